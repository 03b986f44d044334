#!/usr/bin/env bash
# 1 GPU: the whole GPU test suite + smoke + default bench line, as the driver runs them
O=gpurun_out/final_check; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/smoke.log
python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench rc=$? $(python scripts/bench_line.py $O/bench_default.json)"
python bench.py --impl reference > $O/bench_ref_default.json 2> $O/bench_ref_default.err; echo "ref rc=$? $(python scripts/bench_line.py $O/bench_ref_default.json)"
