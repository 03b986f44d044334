"""Tile planner sanity (host-only model)."""
import ctypes as C

import pytest

import wave3d_b200 as w
from wave3d_b200 import _lib
from wave3d_b200.ops import kernels as K
from wave3d_b200.parallel.decomp import Decomposition


@pytest.mark.parametrize("N,nranks,dims", [(512, 1, None), (512, 8, None), (256, 1, None), (1024, 8, (8, 1, 1)),
                                           (2048, 8, None), (33, 1, None), (100, 4, (1, 2, 2))])
def test_plan_is_feasible(N, nranks, dims):
    dec = Decomposition.create(N, nranks, dims)
    for r in {0, nranks - 1}:
        box = dec.box(r)
        pl = K.plan_for(box, 148)
        assert pl.tz == 64 and pl.rj in (1, 2) and 2 <= pl.warps <= 7 and pl.ty == pl.rj * pl.warps
        assert pl.nty * pl.ty >= box.hi[1] - box.lo[1] + 1
        assert pl.ntz * pl.tz >= box.hi[2] - box.lo[2] + 1
        assert pl.smem_bytes <= 227 * 1024 and pl.s1 >= 4 and pl.s0 >= 1
        assert 1 <= pl.ctas <= pl.slots and pl.ctas <= pl.nty * pl.ntz * pl.nchunks
        assert 1 <= pl.nchunks <= _lib.MAX_CHUNKS and pl.xcut[0] == 0 and pl.xcut[pl.nchunks] == box.n[0]
        assert all(pl.xcut[c] < pl.xcut[c + 1] for c in range(pl.nchunks))
        assert sorted(pl.order[:pl.nchunks]) == list(range(pl.nchunks))
        lens = [pl.xcut[pl.order[i] + 1] - pl.xcut[pl.order[i]] for i in range(pl.nchunks)]
        cells = box.n[0] * box.n[1] * box.n[2]
        if cells > 48e6:
            assert lens == sorted(lens, reverse=True)  # one launch per layer, guided: big chunks first in queue order
        else:
            assert max(lens) - min(lens) <= 1          # multi-layer launch: equal chunks
        # the end chunks hold the seam planes (the x-face flags are released when they are done)
        if pl.nchunks > 1:
            assert pl.xcut[1] > box.send_x[0] and pl.xcut[pl.nchunks - 1] <= box.send_x[1]
        assert (0.3 if N >= 256 else 0.0) < pl.efficiency <= 1.0


def test_plan_override(monkeypatch):
    monkeypatch.setenv("WAVE3D_TILE", "12,2,3")
    box = Decomposition.create(256, 1).box(0)
    pl = K.plan_for(box, 148)
    assert (pl.ty, pl.rj, pl.warps, pl.nchunks) == (12, 2, 6, 3)


def test_same_x_extent_means_same_cuts():
    """y/z neighbours share the x extent; the multi-layer style must cut x identically on both sides,
    whatever their y/z extents are (the per-chunk hand-shake of the y/z rank faces relies on it)."""
    dec = Decomposition.create(512, 8, (2, 2, 2))
    plans = {}
    for r in range(8):
        b = dec.box(r)
        pl = _lib.Plan()
        _lib.lib().wave3d_plan_make_style(C.byref(b), 148, 1, C.byref(pl))
        plans.setdefault(b.n[0], []).append(list(pl.xcut[:pl.nchunks + 1]))
    for cuts in plans.values():
        assert all(c == cuts[0] for c in cuts)
