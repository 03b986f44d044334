#!/usr/bin/env bash
# 8 GPUs: multi-GPU tests + parity, strong scaling 4/8, phases at 8 ranks
O=gpurun_out/${1:-s12}; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_parity_reference.py -x -q > $O/pytest_multi.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_multi.log
PORT=29700
tr() { local n=$1 f=$2; shift 2; PORT=$((PORT+1))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $n "$@" > $O/$f.json 2> $O/$f.err
  echo "== $f rc=$? $(python scripts/bench_line.py $O/$f.json)"; }
tr 8 ours_N512_8 --steps 20 --warmup 3
tr 4 ours_N512_4 --steps 20 --warmup 3
WAVE3D_YZ_CHUNK_FLAGS=0 tr 8 ours_N512_8_layerflags --steps 20 --warmup 3 --no-e2e
WAVE3D_CHUNK=16 tr 8 ours_N512_8_c16 --steps 20 --warmup 3 --no-e2e
WAVE3D_CHUNK=64 tr 8 ours_N512_8_c64 --steps 20 --warmup 3 --no-e2e
tr 8 ours_N512_8_1d --steps 20 --warmup 3 --no-e2e --dims 8x1x1
timeout 200 python scripts/cta_phases.py 512 20 1 auto 8 > $O/phases_8gpu.log 2>&1; echo "phases rc=$?"; head -20 $O/phases_8gpu.log | cut -c1-160
