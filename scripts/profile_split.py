#!/usr/bin/env python
"""Eager multi-rank solve on ONE GPU for ncu: profile_split.py N steps dims transport"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import wave3d_b200 as w
N, K = int(sys.argv[1]), int(sys.argv[2])
dims = w.parallel.decomp.parse_dims(sys.argv[3])
tr = sys.argv[4] if len(sys.argv) > 4 else "peer"
p = w.Problem(N=N, T=K * 0.00125, timesteps=K)
nr = dims[0] * dims[1] * dims[2]
with w.WaveSolver(p, nranks=nr, dims=dims, backend="cuda", devices=[0], use_graph=False, transport=tr) as s:
    for _ in range(2):
        r = s.run()
    print(f"N={N} dims={dims} {tr}: {r.solve_ms:.3f} ms {r.gcells_per_s:.1f} Gcell/s plan={s.plan(0).as_dict()}")
