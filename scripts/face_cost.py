#!/usr/bin/env python
"""Cost of the halo push per face type: same N split over 2 ranks along x, y or z.
 - on ONE GPU (ranks time-share it: isolates the instruction cost of the push path)
 - on TWO GPUs in one process (adds NVLink + flag handshake)
usage: face_cost.py [N] [steps] [ngpus]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import wave3d_b200 as w

N = int(sys.argv[1]) if len(sys.argv) > 1 else 512
K = int(sys.argv[2]) if len(sys.argv) > 2 else 20
G = int(sys.argv[3]) if len(sys.argv) > 3 else 1
p = w.Problem(N=N, T=K * 0.00125, timesteps=K)
for nranks, dims in [(1, None), (2, (2, 1, 1)), (2, (1, 2, 1)), (2, (1, 1, 2)), (4, (2, 2, 1)), (8, (2, 2, 2))]:
    if G > 1 and nranks > G:
        continue
    devs = list(range(G)) if G > 1 else [0]
    with w.WaveSolver(p, nranks=nranks, dims=dims, backend="cuda", devices=devs) as s:
        s.run()
        ts = [s.run().solve_ms for _ in range(3)]
        pl = s.plan(0)
        print(f"N={N} ranks={nranks} dims={s.dims} gpus={G}: {min(ts):8.3f} ms  {p.cells*K/min(ts)/1e6:7.1f} Gcell/s  "
              f"plan ty={pl.ty} rj={pl.rj} nchunks={pl.nchunks} ctas={pl.ctas}", flush=True)
