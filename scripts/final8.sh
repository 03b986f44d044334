#!/usr/bin/env bash
# 8 GPUs, final: our arm strong (N=512) + weak (513^3 per GPU) and the 8-rank tests (peer / 1-D / staged / parity)
O=gpurun_out/scale_r2; mkdir -p $O
bash scripts/weak_run.sh 8 0
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_parity_reference.py -v -k "8-" > $O/pytest_8rank.log 2>&1; echo "pytest rc=$?"; tail -9 $O/pytest_8rank.log | cut -c1-150
