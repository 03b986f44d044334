#!/usr/bin/env python
"""Parity of the per-layer maxima against the reference GPU binary (full-precision copy) with several
MPI ranks sharing the visible GPUs. usage: parity_multi.py N nranks [L] [steps]"""
import os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "baseline"))
import run_reference as rr
import wave3d_b200 as w
from wave3d_b200 import _lib
N, P = int(sys.argv[1]), int(sys.argv[2])
Ls = sys.argv[3] if len(sys.argv) > 3 else "1"
K = int(sys.argv[4]) if len(sys.argv) > 4 else 20
L = _lib.PI_FULL if Ls == "pi" else float(Ls)
work = tempfile.mkdtemp()
outs = rr.run_all_ranks_local("cuda_sol", [N, 1, Ls, Ls, Ls, repr(K * 0.00125), K], P, work,
                              binary=os.path.join(rr.REF, "p17", "cuda_sol_p17"),
                              extra_env={"SHIM_CAP_BYTES": str((N + 3) * (N + 3) * 8 + 4096)})
assert all(rc == 0 for rc, _, _ in outs), outs
f = [x for x in os.listdir(work) if x.startswith("output_N")][0]
ms, ra, rrl, _ = rr.parse_output(os.path.join(work, f))
p = w.Problem(N=N, Lx=L, Ly=L, Lz=L, T=K * 0.00125, timesteps=K)
ndev = max(1, _lib.lib().wave3d_cuda_device_count())
with w.WaveSolver(p, nranks=P, backend="cuda", devices=list(range(min(P, ndev)))) as s:
    res = s.run()
with w.WaveSolver(p, nranks=1, backend="cuda") as s:
    one = s.run()
bad_a = [n for n in range(K + 1) if ra[n] != res.max_abs[n]]
bad_r = [n for n in range(K + 1) if rrl[n] != res.max_rel[n]]
print(f"N={N} ranks={P} L={Ls}: abs mismatches {bad_a} rel mismatches {bad_r}; ours P ranks == ours 1 rank: "
      f"{res.max_abs == one.max_abs and res.max_rel == one.max_rel}")
for n in (bad_r[:3] or [K]):
    print(f"  layer {n}: ref {ra[n]!r} {rrl[n]!r} | ours {res.max_abs[n]!r} {res.max_rel[n]!r}")
