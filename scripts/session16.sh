#!/usr/bin/env bash
# 4 GPUs: kernel tests after the k_layer0 / k_field_error rewrite; N=512 on 4 GPUs: launch mode and chunk length
O=gpurun_out/${1:-s16}; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_kernels.py tests/test_zz_no_alloc.py -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
PORT=29900
tr() { local n=$1 f=$2; shift 2; PORT=$((PORT+1))
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $n "$@" > $O/$f.json 2> $O/$f.err
  echo "== $f rc=$? $(python scripts/bench_line.py $O/$f.json)"; }
tr 4 n512_4_default --steps 20 --warmup 3 --no-e2e
WAVE3D_MULTISTEP=0 tr 4 n512_4_perlayer --steps 20 --warmup 3 --no-e2e
WAVE3D_CHUNK=64 tr 4 n512_4_c64 --steps 20 --warmup 3 --no-e2e
WAVE3D_CHUNK=43 tr 4 n512_4_c43 --steps 20 --warmup 3 --no-e2e
WAVE3D_MULTISTEP=1 tr 2 n512_2_multi --steps 20 --warmup 3 --no-e2e
tr 2 n512_2_default --steps 20 --warmup 3 --no-e2e
