#!/usr/bin/env bash
O=gpurun_out/${1:-s10}; mkdir -p $O
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
A256="256 1 1 1 1 0.025 20"
run old256 old/bin/wave3d -- $A256
run c16_s64 bin/wave3d -- $A256
run c32_s64 bin/wave3d WAVE3D_CHUNK=32 -- $A256
run c32_s84 bin/wave3d WAVE3D_CHUNK=32 WAVE3D_TILE=14,2,0,8,4 -- $A256
run c32_s86 bin/wave3d WAVE3D_CHUNK=32 WAVE3D_TILE=14,2,0,8,6 -- $A256
run c32_s66 bin/wave3d WAVE3D_CHUNK=32 WAVE3D_TILE=14,2,0,6,6 -- $A256
run c32_a12 bin/wave3d WAVE3D_CHUNK=32 WAVE3D_AHEAD=12 -- $A256
run c32_nofast bin/wave3d WAVE3D_CHUNK=32 WAVE3D_SCHED=1 -- $A256
run c32_ty7 bin/wave3d WAVE3D_CHUNK=32 WAVE3D_TILE=7,1 -- $A256
run c16_ty7 bin/wave3d WAVE3D_TILE=7,1 -- $A256
run c32_ty12 bin/wave3d WAVE3D_CHUNK=32 WAVE3D_TILE=12,2 -- $A256
run c32_ty10 bin/wave3d WAVE3D_CHUNK=32 WAVE3D_TILE=10,2 -- $A256
run c26 bin/wave3d WAVE3D_CHUNK=26 -- $A256
run c37 bin/wave3d WAVE3D_CHUNK=37 -- $A256
run perlayer bin/wave3d WAVE3D_MULTISTEP=0 -- $A256
WAVE3D_CHUNK=32 timeout 120 python scripts/cta_phases.py 256 20 1 > $O/phases_N256_c32.log 2>&1; tail -15 $O/phases_N256_c32.log
