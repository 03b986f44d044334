#!/usr/bin/env python
"""Small eager (no CUDA graph) solve for ncu captures: python scripts/profile_step.py [N] [steps] [length]."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import wave3d_b200 as w  # noqa: E402
from wave3d_b200 import _lib  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 512
K = int(sys.argv[2]) if len(sys.argv) > 2 else 6
L = sys.argv[3] if len(sys.argv) > 3 else "1"
L = _lib.PI_FULL if L == "pi" else float(L)
p = w.Problem(N=N, Lx=L, Ly=L, Lz=L, T=K * 0.00125, timesteps=K)
with w.WaveSolver(p, nranks=1, backend="cuda", use_graph=False) as s:
    for _ in range(2):
        r = s.run()
    print(f"N={N} K={K} solve {r.solve_ms:.3f} ms  {r.gcells_per_s:.1f} Gcell/s  plan={r.extra['plan']}")
