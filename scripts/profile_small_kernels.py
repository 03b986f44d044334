#!/usr/bin/env python
"""Every kernel besides k_step once, for ncu: tables, layer 0, wall zeroing, face pack / unpack / push, flag
publish, stand-alone error. usage: python scripts/profile_small_kernels.py [N]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import wave3d_b200 as w  # noqa: E402
from wave3d_b200.ops import kernels as K  # noqa: E402
from wave3d_b200.parallel.decomp import Decomposition  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
p = w.Problem(N=N, T=0.005, timesteps=4)
# two ranks on one GPU, eager launches: k_tables, k_layer0, k_face (push of layer 0), k_publish, k_step, k_zero_walls
with w.WaveSolver(p, nranks=2, backend="cuda", devices=[0], use_graph=False) as s:
    r = s.run()
    print("peer solve", r.solve_ms, "ms", r.launches, "launches")
    print("post-pass error", s.field_error(4))   # k_field_error
# staged transport: k_face pack + unpack
with w.WaveSolver(p, nranks=2, backend="cuda", devices=[0], use_graph=False, transport="staged") as s:
    r = s.run()
    print("staged solve", r.solve_ms, "ms", r.launches, "launches")
torch.cuda.synchronize()
