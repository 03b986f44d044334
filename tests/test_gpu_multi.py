"""One rank per GPU under torchrun (NCCL bootstrap): peer transport (cudaIpc-mapped slabs, in-kernel
ghost stores + epoch flags, system-scope atomicMax into rank 0's error table) and the staged NCCL
fallback. Needs as many GPUs as ranks — ranks that spin on each other must never share a GPU."""
import json
import os

import pytest
import torch

import wave3d_b200 as w
from wave3d_b200 import _lib

from test_dist_gloo import launch

pytestmark = pytest.mark.gpu
NGPU = torch.cuda.device_count() if torch.cuda.is_available() else 0


def _ref(N, steps):
    p = w.Problem(N=N, Lx=1.0, Ly=2.0, Lz=_lib.PI_FULL, T=0.05, timesteps=steps)
    with w.WaveSolver(p, nranks=1, backend="cuda") as s:
        return s.run()


@pytest.mark.parametrize("nproc,dims,transport", [(2, "-", "peer"), (2, "1x1x2", "peer"), (2, "-", "staged"),
                                                  (4, "-", "peer"), (8, "-", "peer"), (8, "8x1x1", "peer"),
                                                  (8, "-", "staged")])
def test_one_rank_per_gpu(tmp_path, nproc, dims, transport):
    if NGPU < nproc:
        pytest.skip(f"needs {nproc} GPUs, have {NGPU}")
    out = tmp_path / "res.json"
    r = launch(nproc, ["cuda", transport, 61, 8, dims, out], timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    got = json.load(open(out))
    ref = _ref(61, 8)
    assert got["transport"] == transport
    assert got["abs"] == ref.max_abs and got["rel"] == ref.max_rel
    assert got["abs2"] == ref.max_abs and got["rel2"] == ref.max_rel
    if transport == "peer":
        assert got["extra"].get("root_table_matches_allreduce") is True
        assert got["streamed_ok"] is True


def test_two_processes_share_one_gpu(tmp_path):
    """More processes than GPUs (the reference's MPS launch mode, ref: cuda_sol.cpp:517-519, README.txt:44-46):
    the solver must notice, stage the halo through the host over gloo (no kernel waits on another process's
    flags) and still produce the single-rank maxima bit for bit."""
    if NGPU < 1:
        pytest.skip("needs a GPU")
    out = tmp_path / "res.json"
    r = launch(2, ["cuda", "auto", 45, 6, "-", out], timeout=600, env={"CUDA_VISIBLE_DEVICES": "0"})
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[:3000] + "\n...\n" + r.stderr[-1500:]
    got = json.load(open(out))
    ref = _ref(45, 6)
    assert got["transport"] == "staged"
    assert got["abs"] == ref.max_abs and got["rel"] == ref.max_rel
    assert got["abs2"] == ref.max_abs and got["rel2"] == ref.max_rel
    # asking for the peer transport in this situation is an error that names the alternatives
    r = launch(2, ["cuda", "peer", 45, 6, "-", tmp_path / "x.json"], timeout=600, env={"CUDA_VISIBLE_DEVICES": "0"})
    assert r.returncode != 0 and "one GPU per process" in (r.stdout + r.stderr)
