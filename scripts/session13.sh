#!/usr/bin/env bash
# 2 GPUs: multi-layer launch with rank faces after removing the per-item system fences
O=gpurun_out/${1:-s13}; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_solver.py -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
run() { # run <tag> <exe> <env...> -- args
  local tag=$1 exe=$2; shift 2
  local envs=()
  while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  local best=""
  for i in 1 2; do
    t=$(cd $O && env WAVE3D_QUIET=1 WAVE3D_REPEAT=3 "${envs[@]}" timeout 120 ../../$exe "$@" 2>&1 | grep -o "solve: [0-9.]* ms" | grep -o "[0-9.]*")
    best="$best $t"
  done
  echo "RUN $tag:$best"
}
run n256_1gpu bin/wave3d -- 256 1 1 1 1 0.025 20
run n320_2gpu_y bin/wave3d WAVE3D_DIMS=1x2x1 -- 320 2 1 1 1 0.025 20
run n320_2gpu_z bin/wave3d WAVE3D_DIMS=1x1x2 -- 320 2 1 1 1 0.025 20
run n320_2gpu_x bin/wave3d WAVE3D_DIMS=2x1x1 -- 320 2 1 1 1 0.025 20
run n320_2gpu_y_layerflags bin/wave3d WAVE3D_DIMS=1x2x1 WAVE3D_YZ_CHUNK_FLAGS=0 -- 320 2 1 1 1 0.025 20
run n320_1gpu bin/wave3d -- 320 1 1 1 1 0.025 20
timeout 200 python scripts/cta_phases.py 320 20 1 1x2x1 2 > $O/phases_2gpu_y.log 2>&1; echo "phases rc=$?"; head -17 $O/phases_2gpu_y.log | cut -c1-130
timeout 200 python scripts/cta_phases.py 320 20 1 1x1x2 2 > $O/phases_2gpu_z.log 2>&1; echo "phases rc=$?"; head -17 $O/phases_2gpu_z.log | cut -c1-130
timeout 100 python scripts/nvlink_bytes.py 320 20 2 1x2x1 3 > $O/nvlink_2gpu_y.json 2> $O/nvlink.err; head -30 $O/nvlink_2gpu_y.json
