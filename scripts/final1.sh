#!/usr/bin/env bash
# 1 GPU, final: all GPU tests, both arms at N=512, our arm at N=256 / L=pi
O=gpurun_out/scale_r2; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_1gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest_1gpu.log
bash scripts/weak_run.sh 1 1
for cfg in "256 1" "256 pi" "512 pi"; do
  set -- $cfg
  timeout 200 python bench.py --grid $1 --length $2 --steps 20 --warmup 3 > $O/ours_N$1_L$2_1.json 2> $O/ours_N$1_L$2_1.err; echo "bench $cfg rc=$? $(python scripts/bench_line.py $O/ours_N$1_L$2_1.json)"
done
python -c "import __graft_entry__ as g; g.smoke()"
