#!/usr/bin/env bash
# GPU session 1 (1 GPU): sanity + baseline attribution of the round-1 kernel on small boxes.
O=gpurun_out/s1; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/gpu.txt 2>&1
timeout 300 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"
for cfg in "256 20 1" "256 20 pi" "512 20 1" "128 20 1"; do
  set -- $cfg
  timeout 120 python scripts/cta_phases.py $1 $2 $3 > $O/phases_N$1_L$3.log 2>&1; echo "phases $cfg rc=$?"
done
WAVE3D_MULTISTEP=1 timeout 120 python scripts/cta_phases.py 512 20 1 > $O/phases_N512_multi.log 2>&1
WAVE3D_MULTISTEP=0 timeout 120 python scripts/cta_phases.py 256 20 1 > $O/phases_N256_perlayer.log 2>&1
for L in 1 pi; do
  timeout 200 python bench.py --grid 256 --length $L --steps 20 --warmup 3 --no-e2e > $O/bench_N256_L$L.json 2> $O/bench_N256_L$L.err; echo "bench 256 $L rc=$?"
done
timeout 200 python bench.py --steps 20 --warmup 3 --no-e2e > $O/bench_N512.json 2> $O/bench_N512.err
timeout 200 python bench.py --length pi --steps 20 --warmup 3 --no-e2e > $O/bench_N512_Lpi.json 2> $O/bench_N512_Lpi.err
timeout 100 python scripts/profile_step.py 256 20 1 > $O/prof_plain.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_step -s 3 -c 1 -o $O/prof_multi_N256 python scripts/profile_step.py 256 20 1 > $O/prof_ncu.log 2>&1
echo "ncu rc=$?"
tail -3 $O/pytest.log; cat $O/bench_N256_L1.json $O/bench_N256_Lpi.json | cut -c1-300
