#!/usr/bin/env bash
# Strong (N=512) and weak (513^3 per GPU) scaling lines of both arms on <n> GPUs: scripts/weak_run.sh <n> [both arms strong too: 0|1]
n=${1:-2}; O=gpurun_out/scale_r2; mkdir -p $O
PORT=$((30000 + n * 10))
tr() { local f=$1; shift; PORT=$((PORT+1))
  if [ "$n" = 1 ]; then timeout 900 python bench.py --gpus 1 "$@" > $O/$f.json 2> $O/$f.err
  else timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $n "$@" > $O/$f.json 2> $O/$f.err; fi
  echo "== $f rc=$? $(python scripts/bench_line.py $O/$f.json)"; }
tr ours_strong_$n --steps 20 --warmup 3
[ "$n" != 1 ] && tr ours_weak_$n --scaling weak --steps 20 --warmup 3 --repeats 3
if [ "${2:-0}" = 1 ]; then
  tr ref_strong_$n --impl reference --steps 20 --warmup 3 --repeats 3
  [ "$n" != 1 ] && tr ref_weak_$n --impl reference --scaling weak --steps 20 --warmup 3 --repeats 2
fi
