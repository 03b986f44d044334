#!/usr/bin/env bash
O=gpurun_out/${1:-s8}; mkdir -p $O
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
run new512_alledge bin/wave3d WAVE3D_SCHED=2 -- $A512
run new512_u8 bin/wave3d WAVE3D_TILE=14,2,8 -- $A512
run new512_u16 bin/wave3d WAVE3D_TILE=14,2,16 -- $A512
run new512_u32 bin/wave3d WAVE3D_TILE=14,2,32 -- $A512
run new512_multi_c32 bin/wave3d WAVE3D_MULTISTEP=1 WAVE3D_CHUNK=32 -- $A512
run new512_multi_c64 bin/wave3d WAVE3D_MULTISTEP=1 WAVE3D_CHUNK=64 -- $A512
run new512_multi_c16 bin/wave3d WAVE3D_MULTISTEP=1 -- $A512
run old256 old/bin/wave3d -- $A256
run new256 bin/wave3d -- $A256
run new256_alledge bin/wave3d WAVE3D_SCHED=2 -- $A256
run new256_c32 bin/wave3d WAVE3D_CHUNK=32 -- $A256
run new256_c43 bin/wave3d WAVE3D_CHUNK=43 -- $A256
run new256_c64 bin/wave3d WAVE3D_CHUNK=64 -- $A256
run new256pi_c32 bin/wave3d WAVE3D_CHUNK=32 -- 256 1 pi pi pi 0.025 20
run old256pi old/bin/wave3d -- 256 1 pi pi pi 0.025 20
