#!/usr/bin/env bash
# Scaling sweep on one multi-GPU box. Usage: scripts/scale_run.sh <max_gpus> <tag> [ours|ref|all]
# strong scaling N=512 at 1/2/4/8, N=1024 3-D vs 1-D grid, N=2048 on 8 GPUs (weak partner of N=1024 on 1).
G=${1:-8}; TAG=${2:-r1}; WHAT=${3:-all}; OUT=gpurun_out/scale_$TAG; mkdir -p $OUT
PORT=29600
tr() { # tr <nproc> <outfile> <bench args...>
  local n=$1 f=$2; shift 2; PORT=$((PORT+1))
  if [ "$n" = 1 ]; then timeout 600 python bench.py --gpus 1 "$@" > $OUT/$f.json 2> $OUT/$f.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $n "$@" > $OUT/$f.json 2> $OUT/$f.err; fi
  echo "== $f rc=$? $(python scripts/bench_line.py $OUT/$f.json)"
}
nvidia-smi topo -m > $OUT/topo.txt 2>&1
if [ $WHAT != ref ]; then
  for n in 1 2 4 8; do [ $n -le $G ] && tr $n ours_N512_$n --steps 20 --warmup 3; done
  if [ $G -ge 8 ]; then
    tr 8 ours_N1024_8_3d --grid 1024 --steps 20 --warmup 3
    tr 8 ours_N1024_8_1d --grid 1024 --steps 20 --warmup 3 --dims 8x1x1
    tr 8 ours_N512_8_staged --steps 20 --warmup 3 --transport staged
    tr 8 ours_N2048_8 --grid 2048 --steps 10 --warmup 3 --repeats 3
  fi
  tr 1 ours_N1024_1 --grid 1024 --steps 20 --warmup 3 --repeats 3
  tr 1 ours_N256pi_1 --grid 256 --length pi --steps 20 --warmup 3
fi
if [ $WHAT != ours ]; then
  for n in 1 2 4 8; do [ $n -le $G ] && tr $n ref_N512_$n --impl reference --steps 20 --warmup 3; done
  if [ $G -ge 8 ]; then
    tr 8 ref_N1024_8 --impl reference --grid 1024 --steps 20 --warmup 3 --repeats 2
    tr 8 ref_N2048_8 --impl reference --grid 2048 --steps 10 --warmup 3 --repeats 2
  fi
  tr 1 ref_N256pi_1 --impl reference --grid 256 --length pi --steps 20 --warmup 3
fi
