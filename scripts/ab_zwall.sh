#!/usr/bin/env bash
# 1 GPU: GPU tests, then warm A/B of "z+ wall tiles take the edge variant" (WAVE3D_SCHED=4, the behaviour before) vs the default
O=gpurun_out/${1:-ab_zwall}; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest.log
run() { local tag=$1 sched=$2; shift 2; local best=""
  for i in 1 2 3; do
    t=$(cd $O && env WAVE3D_QUIET=1 WAVE3D_PHASES=0 WAVE3D_REPEAT=4 WAVE3D_SCHED=$sched timeout 120 ../../bin/wave3d "$@" 2>&1 | grep -o "solve: [0-9.]* ms" | tail -1 | grep -o "[0-9.]*")
    best="$best $t"
  done
  echo "RUN $tag:$best"; }
for cfg in "256 1 1 1 1" "256 1 pi pi pi" "512 1 1 1 1" "512 1 pi pi pi" "384 1 1 1 1" "128 1 1 1 1"; do
  set -- $cfg
  run edge_N$1_L$3 4 $cfg 0.025 20
  run wall_N$1_L$3 0 $cfg 0.025 20
done
python bench.py --grid 256 > $O/bench_256.json 2>/dev/null; python scripts/bench_line.py $O/bench_256.json
python bench.py --grid 256 --length pi > $O/bench_256pi.json 2>/dev/null; python scripts/bench_line.py $O/bench_256pi.json
python bench.py > $O/bench_512.json 2>/dev/null; python scripts/bench_line.py $O/bench_512.json
