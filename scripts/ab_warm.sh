#!/usr/bin/env bash
# 1 GPU: GPU tests, then WARM A/B of the round-1 binary (old/, built from d6241b0 + a WAVE3D_REPEAT loop) vs the current one
O=gpurun_out/${1:-ab}; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest.log
run() { local tag=$1 exe=$2; shift 2; local best=""
  for i in 1 2 3; do
    t=$(cd $O && env WAVE3D_QUIET=1 WAVE3D_REPEAT=4 timeout 120 ../../$exe "$@" 2>&1 | grep -o "solve: [0-9.]* ms" | grep -o "[0-9.]*")
    best="$best $t"
  done
  echo "RUN $tag:$best"; }
for cfg in "512 1 1 1 1" "512 1 pi pi pi" "256 1 1 1 1" "256 1 pi pi pi" "384 1 1 1 1" "1024 1 1 1 1"; do
  set -- $cfg
  [ -x old/bin/wave3d ] && run old_N$1_L$3 old/bin/wave3d $cfg 0.025 20
  run new_N$1_L$3 bin/wave3d $cfg 0.025 20
done
timeout 120 python scripts/cta_phases.py 512 20 1 > $O/phases_N512.log 2>&1; tail -6 $O/phases_N512.log
