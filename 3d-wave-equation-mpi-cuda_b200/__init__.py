"""wave3d-b200 — B200-native 3-D wave-equation solver (Python face of libwave3d).

The directory name follows the project naming scheme and is not a valid Python
identifier; import the package through its alias ``wave3d_b200`` (repo root), which
puts this directory on its ``__path__``.

Layout
  models/    WaveSolver — the flagship solve (single process, 1..8 GPUs or CPU) and
             DistributedWaveSolver — one rank per process under torch.distributed
  ops/       kernel-level entry points on torch tensors + plain-PyTorch fp64 references
  parallel/  block decomposition, peer wiring over torch.distributed, staged NCCL/gloo halo
  utils/     problem/CLI spec, report writer, library loader/builder, clock sampler
"""
