#!/usr/bin/env python
"""Where a layer's time goes inside the fused step kernel: per-CTA phase accumulators (WAVE3D_DBG_TS=1).

usage: python scripts/cta_phases.py N steps [length] [dims] [ngpus]
Runs in ONE process (ranks on `ngpus` devices, peer transport) and prints, per rank, the median / max over
CTAs of every phase as a fraction of the kernel's cycles: producer lane (queue fetch, dependency poll,
acquire fences, waiting for a free ring slot = ahead of the compute warps, publishing) and compute warp 0
(waiting for an item id, for the first planes of an item, for planes inside the march, set-up, march)."""
import ctypes as C
import os
import sys

os.environ["WAVE3D_DBG_TS"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import wave3d_b200 as w  # noqa: E402
from wave3d_b200 import _lib  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
K = int(sys.argv[2]) if len(sys.argv) > 2 else 20
L = sys.argv[3] if len(sys.argv) > 3 else "1"
L = _lib.PI_FULL if L == "pi" else float(L)
dims = w.parallel.decomp.parse_dims(sys.argv[4]) if len(sys.argv) > 4 and sys.argv[4] != "auto" else None
ng = int(sys.argv[5]) if len(sys.argv) > 5 else 1
p = w.Problem(N=N, Lx=L, Ly=L, Lz=L, T=K * 0.00125, timesteps=K)
names = ["queue", "poll", "acquire", "slot_wait", "publish", "items", "items_waited", "poll_face", "face_items",
         "kernel_cyc", "c:item_wait", "c:first_planes", "c:plane_wait", "c:setup", "c:march"]
with w.WaveSolver(p, nranks=ng, dims=dims, backend="cuda", devices=list(range(ng)), use_graph=False) as s:
    s.run()
    r = s.run()
    print(f"N={N} K={K} ranks={ng} dims={r.dims} solve {r.solve_ms:.3f} ms = {r.solve_ms / K * 1e3:.1f} us/step "
          f"{r.gcells_per_s:.1f} Gcell/s multi={s.multistep_active} plan={s.plan(0).as_dict()}")
    lib = _lib.lib()
    lib.wave3d_debug_copy_phases.argtypes = [C.c_void_p, C.c_int, C.c_int]
    for g in range(ng):
        n = s.plan(g).ctas
        buf = (C.c_ulonglong * (16 * n))()
        assert lib.wave3d_debug_copy_phases(buf, n, g) == 0
        a = np.frombuffer(buf, dtype=np.uint64).reshape(n, 16).astype(np.float64)
        kc = a[:, 9]
        print(f" rank/gpu {g}: ctas={n} kernel cycles median {np.median(kc):.0f} (last launch)")
        for i, nm in enumerate(names):
            if nm in ("items", "items_waited", "face_items"):
                print(f"   {nm:14s} sum {a[:, i].sum():9.0f}  per CTA median {np.median(a[:, i]):7.1f} max {a[:, i].max():7.0f}")
            elif nm != "kernel_cyc":
                f = a[:, i] / np.maximum(kc, 1)
                print(f"   {nm:14s} median {100 * np.median(f):6.2f}%  max {100 * f.max():6.2f}%  "
                      f"(mean {a[:, i].mean() / max(1.0, a[:, 5].mean()):8.0f} cyc/item)")
