#!/usr/bin/env bash
# 1 GPU: ncu --set full of the final step kernel (N=512 one launch per layer, N=256 multi-layer launch L=1 and L=pi)
O=gpurun_out/final_ncu; mkdir -p $O
timeout 100 python scripts/profile_step.py 512 4 1 > $O/p512_plain.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_step -s 5 -c 1 -o $O/prof_N512 python scripts/profile_step.py 512 4 1 > $O/p512_ncu.log 2>&1
echo "ncu 512 rc=$?"
timeout 100 python scripts/profile_step.py 256 20 1 > $O/p256_plain.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_step -s 3 -c 1 -o $O/prof_multi_N256 python scripts/profile_step.py 256 20 1 > $O/p256_ncu.log 2>&1
echo "ncu 256 rc=$?"
timeout 100 python scripts/profile_step.py 256 20 pi > $O/p256pi_plain.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_step -s 3 -c 1 -o $O/prof_multi_N256_pi python scripts/profile_step.py 256 20 pi > $O/p256pi_ncu.log 2>&1
echo "ncu 256pi rc=$?"
