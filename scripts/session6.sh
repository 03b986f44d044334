#!/usr/bin/env bash
O=gpurun_out/${1:-s6}; mkdir -p $O
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
run new512_a0 bin/wave3d WAVE3D_AHEAD=0 -- $A512
run new512_a1 bin/wave3d WAVE3D_AHEAD=1 -- $A512
run new512_a6 bin/wave3d WAVE3D_AHEAD=6 -- $A512
run new512_nofast bin/wave3d WAVE3D_SCHED=1 -- $A512
run new512_nofast_a0 bin/wave3d WAVE3D_SCHED=1 WAVE3D_AHEAD=0 -- $A512
run old256 old/bin/wave3d -- $A256
run new256 bin/wave3d -- $A256
run new256_nofast bin/wave3d WAVE3D_SCHED=1 -- $A256
run new256_a3 bin/wave3d WAVE3D_AHEAD=3 -- $A256
run new256_a10 bin/wave3d WAVE3D_AHEAD=10 -- $A256
run new256_c32 bin/wave3d WAVE3D_CHUNK=32 -- $A256
python scripts/cta_timeline.py 512 6 > $O/timeline_N512.log 2>&1; cat $O/timeline_N512.log
