#!/usr/bin/env bash
# 8 GPUs (lean): strong scaling N=512 at 8 and 4 after removing per-item system fences; NVLink counters
O=gpurun_out/${1:-s14}; mkdir -p $O
PORT=29800
tr() { local n=$1 f=$2; shift 2; PORT=$((PORT+1))
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $n "$@" > $O/$f.json 2> $O/$f.err
  echo "== $f rc=$? $(python scripts/bench_line.py $O/$f.json)"; }
nvidia-smi nvlink -gt d > $O/nvlink_before.txt 2>&1
tr 8 ours_N512_8 --steps 20 --warmup 3
nvidia-smi nvlink -gt d > $O/nvlink_after.txt 2>&1
tr 4 ours_N512_4 --steps 20 --warmup 3 --no-e2e
timeout 100 python scripts/cta_phases.py 512 20 1 auto 8 > $O/phases_8gpu.log 2>&1; echo "phases rc=$?"; head -2 $O/phases_8gpu.log | cut -c1-100
head -12 $O/nvlink_before.txt; head -12 $O/nvlink_after.txt
