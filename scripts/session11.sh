#!/usr/bin/env bash
# 2 GPUs: multi-GPU correctness with the new kernel (per-chunk y/z flags) + parity + bench lines
O=gpurun_out/${1:-s11}; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_parity_reference.py tests/test_gpu_solver.py -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log
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
run n256_1gpu_multi bin/wave3d -- 256 1 1 1 1 0.025 20
run n256_1gpu_perlayer bin/wave3d WAVE3D_MULTISTEP=0 -- 256 1 1 1 1 0.025 20
run n512_1gpu bin/wave3d -- 512 1 1 1 1 0.025 20
run n512_2gpu_x bin/wave3d -- 512 2 1 1 1 0.025 20
run n512_2gpu_y bin/wave3d WAVE3D_DIMS=1x2x1 -- 512 2 1 1 1 0.025 20
run n512_2gpu_z bin/wave3d WAVE3D_DIMS=1x1x2 -- 512 2 1 1 1 0.025 20
run n512_2gpu_y_layerflags bin/wave3d WAVE3D_DIMS=1x2x1 WAVE3D_YZ_CHUNK_FLAGS=0 -- 512 2 1 1 1 0.025 20
run n512_2gpu_z_layerflags bin/wave3d WAVE3D_DIMS=1x1x2 WAVE3D_YZ_CHUNK_FLAGS=0 -- 512 2 1 1 1 0.025 20
run n256x512x512_y bin/wave3d WAVE3D_DIMS=1x2x1 -- 320 2 1 1 1 0.025 20
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 3 > $O/bench_2gpu.json 2> $O/bench_2gpu.err; echo "bench2 rc=$? $(python scripts/bench_line.py $O/bench_2gpu.json)"
python bench.py --gpus 1 --steps 20 --warmup 3 > $O/bench_1gpu.json 2> $O/bench_1gpu.err; echo "bench1 rc=$? $(python scripts/bench_line.py $O/bench_1gpu.json)"
python bench.py --gpus 1 --grid 256 --steps 20 --warmup 3 > $O/bench_1gpu_256.json 2> $O/bench_1gpu_256.err; echo "bench1_256 rc=$? $(python scripts/bench_line.py $O/bench_1gpu_256.json)"
