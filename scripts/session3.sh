#!/usr/bin/env bash
# GPU session (1 GPU): tests, same-box A/B of the round-1 binary (old/) vs the current one, phases, bench lines
O=gpurun_out/${1:-s3}; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
ab() { # ab <tag> <args...>
  local tag=$1; shift
  for v in old new; do
    local exe=bin/wave3d; [ $v = old ] && exe=old/bin/wave3d
    local best=""
    for i in 1 2 3; do
      t=$(cd $O && WAVE3D_QUIET=1 timeout 120 ../../$exe "$@" 2>&1 | grep -o "solve: [0-9.]* ms" | grep -o "[0-9.]*")
      best="$best $t"
    done
    echo "AB $tag $v:$best"
  done
}
ab N512 512 1 1 1 1 0.025 20
ab N256 256 1 1 1 1 0.025 20
ab N256pi 256 1 pi pi pi 0.025 20
ab N512pi 512 1 pi pi pi 0.025 20
for cfg in "256 20 1" "512 20 1"; do
  set -- $cfg
  timeout 120 python scripts/cta_phases.py $1 $2 $3 > $O/phases_N$1_L$3.log 2>&1; echo "phases $cfg rc=$?"; head -1 $O/phases_N$1_L$3.log | cut -c1-120
done
WAVE3D_MULTISTEP=1 timeout 120 python scripts/cta_phases.py 512 20 1 > $O/phases_N512_multi.log 2>&1; head -1 $O/phases_N512_multi.log | cut -c1-120
for cfg in "256 1" "256 pi" "512 1" "512 pi"; do
  set -- $cfg
  timeout 200 python bench.py --grid $1 --length $2 --steps 20 --warmup 3 --no-e2e > $O/bench_N$1_L$2.json 2> $O/bench_N$1_L$2.err; echo "bench $cfg rc=$? $(python scripts/bench_line.py $O/bench_N$1_L$2.json)"
done
timeout 100 python scripts/profile_step.py 256 20 1 > $O/prof_plain.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_step -s 3 -c 1 -o $O/prof_multi_N256 python scripts/profile_step.py 256 20 1 > $O/prof_ncu.log 2>&1
echo "ncu rc=$?"
