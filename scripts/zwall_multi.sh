#!/usr/bin/env bash
# G GPUs (2 or 4): the <=G-rank multi-GPU and parity tests, then the strong-scaling points of bench.py on this box
G=${1:-2}; O=gpurun_out/zwall_multi_$G; mkdir -p $O
K="2- or -2]"; [ $G -ge 4 ] && K="4- or -2]"
timeout 240 python -m pytest tests/test_gpu_multi.py tests/test_gpu_parity_reference.py -v -k "$K" > $O/pytest.log 2>&1; echo "pytest rc=$?"; grep -E "PASSED|FAILED|ERROR|passed|failed" $O/pytest.log | cut -c1-160
PORT=29700
for n in 1 2 4; do
  [ $n -le $G ] || continue
  [ $G -ge 4 ] && [ $n -lt 4 ] && continue
  PORT=$((PORT+1)); f=ours_N512_$n
  if [ $n = 1 ]; then timeout 300 python bench.py --steps 20 --warmup 3 > $O/$f.json 2> $O/$f.err
  else timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $n --steps 20 --warmup 3 > $O/$f.json 2> $O/$f.err; fi
  echo "== $f rc=$? $(python scripts/bench_line.py $O/$f.json)"
done
