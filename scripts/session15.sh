#!/usr/bin/env bash
# 1 GPU: tests, launch list of a whole N=256 solve, ncu digests (small kernels, N=256 multi L=1 and L=pi)
O=gpurun_out/${1:-s15}; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
timeout 100 python scripts/profile_step.py 256 20 1 > $O/p256_plain.log 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file $O/launches_N256.csv python scripts/profile_step.py 256 20 1 > $O/p256_ncu.log 2>&1
echo "launch list rc=$?"
timeout 100 python scripts/profile_step.py 512 20 1 > $O/p512_plain.log 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/launches_N512.csv python scripts/profile_step.py 512 20 1 > $O/p512_ncu.log 2>&1
echo "launch list 512 rc=$?"
timeout 100 python scripts/profile_small_kernels.py 256 > $O/small_plain.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"k_tables|k_layer0|k_face|k_zero_walls|k_field_error|k_publish" -c 14 -o $O/prof_small python scripts/profile_small_kernels.py 256 > $O/small_ncu.log 2>&1
echo "ncu small rc=$?"
timeout 100 python scripts/profile_step.py 256 20 pi > $O/p256pi_plain.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_step -s 3 -c 1 -o $O/prof_multi_N256_pi python scripts/profile_step.py 256 20 pi > $O/p256pi_ncu.log 2>&1
echo "ncu 256pi rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_step -s 3 -c 1 -o $O/prof_multi_N256 python scripts/profile_step.py 256 20 1 > $O/p256_ncu2.log 2>&1
echo "ncu 256 rc=$?"
for cfg in "256 1" "256 pi" "512 1" "512 pi"; do
  set -- $cfg
  timeout 200 python bench.py --grid $1 --length $2 --steps 20 --warmup 3 > $O/bench_N$1_L$2.json 2> $O/bench_N$1_L$2.err; echo "bench $cfg rc=$? $(python scripts/bench_line.py $O/bench_N$1_L$2.json)"
done
