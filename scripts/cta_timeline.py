#!/usr/bin/env python
"""Per-CTA phase timeline of the last fused-step launch (debug build switch WAVE3D_DBG_TS=1).
usage: WAVE3D_DBG_TS=1 python scripts/cta_timeline.py [N] [steps]"""
import ctypes as C, os, sys
os.environ["WAVE3D_DBG_TS"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import wave3d_b200 as w
from wave3d_b200 import _lib
N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
K = int(sys.argv[2]) if len(sys.argv) > 2 else 8
p = w.Problem(N=N, T=K * 0.00125, timesteps=K)
with w.WaveSolver(p, backend="cuda", use_graph=False) as s:
    s.run(); r = s.run()
    n = s.plan(0).ctas
    buf = (C.c_ulonglong * (8 * n))()
    lib = _lib.lib(); lib.wave3d_debug_copy_timestamps.argtypes = [C.c_void_p, C.c_int]
    assert lib.wave3d_debug_copy_timestamps(buf, n) == 0
    t = np.frombuffer(buf, dtype=np.uint64).reshape(n, 8).astype(np.float64)
    t0 = t[:, 0].min()
    names = ["enter", "init_done", "first_data", "loop_end", "after_sync", "after_fence"]
    print(f"N={N} solve {r.solve_ms:.3f} ms, {r.solve_ms/K*1e3:.1f} us/step, ctas={n}; times in us relative to the first CTA entry")
    for i, nm in enumerate(names):
        col = (t[:, i] - t0) / 1e3
        print(f"  {nm:12s} min {col.min():8.2f}  median {np.median(col):8.2f}  max {col.max():8.2f}")
