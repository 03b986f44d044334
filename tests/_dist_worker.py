"""Worker for the multi-process tests: one rank of a DistributedWaveSolver run (torchrun env)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import wave3d_b200 as w  # noqa: E402
from wave3d_b200 import _lib  # noqa: E402
from wave3d_b200.parallel import dist as pdist  # noqa: E402

backend = sys.argv[1]            # cpu | cuda
transport = sys.argv[2]          # staged | peer | auto
N, steps = int(sys.argv[3]), int(sys.argv[4])
dims = None if sys.argv[5] == "-" else w.parallel.decomp.parse_dims(sys.argv[5])
out = sys.argv[6]
pdist.init_process_group_from_env("gloo" if backend == "cpu" else "nccl")
p = w.Problem(N=N, Lx=1.0, Ly=2.0, Lz=_lib.PI_FULL, T=0.05, timesteps=steps)
s = w.DistributedWaveSolver(p, dims=dims, backend=backend, transport=transport)
r1 = s.run()
r2 = s.run()  # a second solve on the same solver: reset / flag epochs must be reusable
streamed_ok = None
if backend == "cuda" and s.transport == "peer":
    # end-to-end flavour on every rank: per-layer uploads and read-backs. Each rank reads its own maxima
    # (rank 0's table may already hold forwarded ones); their maximum over ranks must equal the batch solve's.
    s._reset()
    rs = s.inner.run_streamed()
    s._barrier()
    red = pdist.max_over_ranks(rs.max_abs + rs.max_rel, s.coll_device)
    L = p.timesteps + 1
    streamed_ok = red[:L] == r2.max_abs and red[L:] == r2.max_rel
if dist.get_rank() == 0:
    json.dump({"abs": r1.max_abs, "rel": r1.max_rel, "abs2": r2.max_abs, "rel2": r2.max_rel, "dims": list(r1.dims),
               "transport": s.transport, "ms": r2.solve_ms, "extra": {k: v for k, v in r2.extra.items() if k != "plan"},
               "launches": r2.launches, "streamed_ok": streamed_ok}, open(out, "w"))
    s.write_report(r2, out + ".txt")
s.close()
dist.destroy_process_group()
