#!/usr/bin/env bash
# GPU session (1 GPU): correctness of the rewritten step kernel + phase attribution + bench lines
O=gpurun_out/${1:-s2}; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log
for cfg in "256 20 1" "256 20 pi" "512 20 1"; do
  set -- $cfg
  timeout 120 python scripts/cta_phases.py $1 $2 $3 > $O/phases_N$1_L$3.log 2>&1; echo "phases $cfg rc=$?"; head -1 $O/phases_N$1_L$3.log | cut -c1-120
done
WAVE3D_MULTISTEP=1 timeout 120 python scripts/cta_phases.py 512 20 1 > $O/phases_N512_multi.log 2>&1; head -1 $O/phases_N512_multi.log | cut -c1-120
for cfg in "256 1" "256 pi" "512 1" "512 pi"; do
  set -- $cfg
  timeout 200 python bench.py --grid $1 --length $2 --steps 20 --warmup 3 > $O/bench_N$1_L$2.json 2> $O/bench_N$1_L$2.err; echo "bench $cfg rc=$? $(python scripts/bench_line.py $O/bench_N$1_L$2.json)"
done
