#!/usr/bin/env bash
O=gpurun_out/${1:-s7}; mkdir -p $O
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
run new512_s54 bin/wave3d WAVE3D_TILE=14,2,0,5,4 -- $A512
run new512_s74 bin/wave3d WAVE3D_TILE=14,2,0,7,4 -- $A512
run new512_s63 bin/wave3d WAVE3D_TILE=14,2,0,6,3 -- $A512
run new512_s65 bin/wave3d WAVE3D_TILE=14,2,0,6,5 -- $A512
run old256 old/bin/wave3d -- $A256
run new256 bin/wave3d -- $A256
run new256_c32 bin/wave3d WAVE3D_CHUNK=32 -- $A256
run new256_c24 bin/wave3d WAVE3D_CHUNK=24 -- $A256
run new256_s74 bin/wave3d WAVE3D_TILE=14,2,0,7,4 -- $A256
run old256pi old/bin/wave3d -- 256 1 pi pi pi 0.025 20
run new256pi bin/wave3d -- 256 1 pi pi pi 0.025 20
run old512pi old/bin/wave3d -- 512 1 pi pi pi 0.025 20
run new512pi bin/wave3d -- 512 1 pi pi pi 0.025 20
for cfg in "256 20 1" "512 20 1"; do
  set -- $cfg
  timeout 120 python scripts/cta_phases.py $1 $2 $3 > $O/phases_N$1_L$3.log 2>&1; echo "phases $cfg rc=$?"; head -1 $O/phases_N$1_L$3.log | cut -c1-120
done
timeout 100 python scripts/profile_step.py 512 4 1 > $O/prof_plain.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_step -s 5 -c 1 -o $O/prof_N512 python scripts/profile_step.py 512 4 1 > $O/prof_ncu.log 2>&1
echo "ncu rc=$?"
