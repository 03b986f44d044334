"""One rank per process over torch.distributed (gloo, CPU backend, staged halo): the host-side logic of
the multi-GPU product path — rank/neighbour wiring, message ordering, reduction of the maxima, report."""
import json
import os
import socket
import subprocess
import sys

import pytest

import wave3d_b200 as w
from wave3d_b200 import _lib

WORKER = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_dist_worker.py")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def launch(nproc, args, timeout=300, env=None):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), WORKER] + [str(a) for a in args]
    env = dict(os.environ, OMP_NUM_THREADS="1", **(env or {}))
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env)


@pytest.mark.parametrize("nproc,dims", [(2, "-"), (4, "1x2x2"), (3, "3x1x1")])
def test_gloo_staged_matches_single_process(tmp_path, nproc, dims):
    out = tmp_path / "res.json"
    r = launch(nproc, ["cpu", "staged", 29, 6, dims, out])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    got = json.load(open(out))
    p = w.Problem(N=29, Lx=1.0, Ly=2.0, Lz=_lib.PI_FULL, T=0.05, timesteps=6)
    with w.WaveSolver(p, nranks=1, backend="cpu") as s:
        ref = s.run()
    assert got["abs"] == ref.max_abs and got["rel"] == ref.max_rel
    assert got["abs2"] == ref.max_abs and got["rel2"] == ref.max_rel
    assert got["transport"] == "staged"
    txt = open(str(out) + ".txt").read().splitlines()
    assert txt[0].startswith("numerical solution calculated in ") and len(txt) == 1 + 7
