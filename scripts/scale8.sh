#!/usr/bin/env bash
# Lean 8-GPU sweep of the new solver only (reference numbers: scripts/scale_run.sh ... ref). Usage: scale8.sh <tag>
TAG=${1:-r1c}; OUT=gpurun_out/scale_$TAG; mkdir -p $OUT; PORT=29800
tr() { local n=$1 f=$2; shift 2; PORT=$((PORT+1))
  if [ "$n" = 1 ]; then timeout 600 python bench.py --gpus 1 "$@" > $OUT/$f.json 2> $OUT/$f.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $n "$@" > $OUT/$f.json 2> $OUT/$f.err; fi
  echo "== $f rc=$? $(python scripts/bench_line.py $OUT/$f.json)"; }
for n in 1 2 4 8; do tr $n ours_N512_$n --steps 20 --warmup 3; done
WAVE3D_MULTISTEP=0 tr 8 ours_N512_8_perstep --steps 20 --warmup 3 --no-e2e
WAVE3D_MULTISTEP=0 tr 4 ours_N512_4_perstep --steps 20 --warmup 3 --no-e2e
tr 8 ours_N1024_8_3d --grid 1024 --steps 20 --warmup 3
tr 8 ours_N1024_8_1d --grid 1024 --steps 20 --warmup 3 --dims 8x1x1
tr 8 ours_N512_8_staged --steps 20 --warmup 3 --transport staged --no-e2e
tr 8 ours_N2048_8 --grid 2048 --steps 10 --warmup 3 --repeats 3
