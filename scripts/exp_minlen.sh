#!/usr/bin/env bash
# 1 GPU: guided chunk floor (WAVE3D_MINLEN) at N=512 / N=1024, one launch per layer, warm
O=gpurun_out/${1:-exp_minlen}; mkdir -p $O
run() { local tag=$1; shift; local envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  local best=""
  for i in 1 2; do
    t=$(cd $O && env WAVE3D_QUIET=1 WAVE3D_REPEAT=4 "${envs[@]}" timeout 120 ../../bin/wave3d "$@" 2>&1 | grep -o "solve: [0-9.]* ms" | grep -o "[0-9.]*")
    best="$best $t"
  done
  echo "RUN $tag:$best"; }
for m in 8 16 24 32 48 64 96; do run n512_minlen$m WAVE3D_MINLEN=$m -- 512 1 1 1 1 0.025 20; done
for m in 32 64 128; do run n1024_minlen$m WAVE3D_MINLEN=$m -- 1024 1 1 1 1 0.025 20; done
for t in "14,2,3" "14,2,4" "14,2,5"; do run n512_tile$t WAVE3D_TILE=$t -- 512 1 1 1 1 0.025 20; done
