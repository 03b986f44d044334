#!/usr/bin/env bash
TAG=${1:-r1d}; OUT=gpurun_out/scale_$TAG; mkdir -p $OUT; PORT=29900
tr() { local n=$1 f=$2; shift 2; PORT=$((PORT+1))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $n "$@" > $OUT/$f.json 2> $OUT/$f.err
  echo "== $f rc=$? $(python scripts/bench_line.py $OUT/$f.json)"; }
tr 8 ours_N512_8 --steps 20 --warmup 3
tr 8 ours_N512_8_1d --steps 20 --warmup 3 --dims 8x1x1 --no-e2e
tr 4 ours_N512_4 --steps 20 --warmup 3 --no-e2e
