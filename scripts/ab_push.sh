#!/usr/bin/env bash
# G GPUs: A/B of "full tiles on a y/z rank face take the general march" (WAVE3D_SCHED=8, before) vs MODE 2 (default)
G=${1:-2}; DIMS=${2:-1x1x2}; O=gpurun_out/ab_push_$G; mkdir -p $O; PORT=29800
for rep in 1 2; do for s in 8 0; do
  PORT=$((PORT+1)); f=N512_${DIMS}_sched${s}_$rep
  WAVE3D_SCHED=$s timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $G --steps 20 --warmup 3 --dims $DIMS --no-e2e > $O/$f.json 2> $O/$f.err
  echo "== $f rc=$? $(python scripts/bench_line.py $O/$f.json)"
done; done
