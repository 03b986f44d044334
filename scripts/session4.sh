#!/usr/bin/env bash
O=gpurun_out/${1:-s4}; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.log
run() { # run <tag> <exe> <env...> -- args
  local tag=$1 exe=$2; shift 2
  local envs=()
  while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  local best=""
  for i in 1 2 3; do
    t=$(cd $O && env WAVE3D_QUIET=1 "${envs[@]}" timeout 120 ../../$exe "$@" 2>&1 | grep -o "solve: [0-9.]* ms" | grep -o "[0-9.]*")
    best="$best $t"
  done
  echo "RUN $tag:$best"
}
A512="512 1 1 1 1 0.025 20"; A256="256 1 1 1 1 0.025 20"
run old512 old/bin/wave3d -- $A512
run new512 bin/wave3d -- $A512
run new512_s84 bin/wave3d WAVE3D_TILE=14,2,0,8,4 -- $A512
run new512_s44 bin/wave3d WAVE3D_TILE=14,2,0,4,4 -- $A512
run new512_s88 bin/wave3d WAVE3D_TILE=14,2,0,8,8 -- $A512
run new512_s42 bin/wave3d WAVE3D_TILE=14,2,0,4,2 -- $A512
run old256 old/bin/wave3d -- $A256
run new256 bin/wave3d -- $A256
run new256_c8 bin/wave3d WAVE3D_CHUNK=8 -- $A256
run new256_c12 bin/wave3d WAVE3D_CHUNK=12 -- $A256
run new256_c24 bin/wave3d WAVE3D_CHUNK=24 -- $A256
run new256_c32 bin/wave3d WAVE3D_CHUNK=32 -- $A256
run new256_s44 bin/wave3d WAVE3D_TILE=14,2,0,4,4 -- $A256
run new512multi bin/wave3d WAVE3D_MULTISTEP=1 -- $A512
run old256pi old/bin/wave3d -- 256 1 pi pi pi 0.025 20
run new256pi bin/wave3d -- 256 1 pi pi pi 0.025 20
for cfg in "256 20 1" "512 20 1"; do
  set -- $cfg
  timeout 120 python scripts/cta_phases.py $1 $2 $3 > $O/phases_N$1_L$3.log 2>&1; echo "phases $cfg rc=$?"; head -1 $O/phases_N$1_L$3.log | cut -c1-120
done
