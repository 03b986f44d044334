"""Numerics of every sm_100a kernel against the plain-PyTorch fp64 reference of the same op and
(bit-exactly) against the C++ CPU twin, on small odd-shaped boxes covering all 27 boundary /
neighbour combinations (SURVEY §4.2 'unit (GPU)')."""
import ctypes as C

import pytest
import torch

import wave3d_b200 as w
from wave3d_b200 import _lib
from wave3d_b200.ops import kernels as K, reference as R
from wave3d_b200.parallel.decomp import Decomposition

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _ulps(a: torch.Tensor, b: torch.Tensor) -> int:
    ia, ib = a.contiguous().view(torch.int64), b.contiguous().view(torch.int64)
    return int((ia - ib).abs().max())


def test_native_library_loaded_and_device_visible():
    assert _lib.cuda_available(), "libwave3d sees no CUDA device"
    assert torch.cuda.is_available()


def test_tables_match_reference_expressions():
    p = w.Problem(N=97, Lx=1.0, Ly=2.0, Lz=_lib.PI_FULL, T=0.1, timesteps=4)
    dec = Decomposition.create(p.N, 8)
    for r in (0, 7):
        box = dec.box(r)
        got = K.tables(p, box, DEV)
        want = R.tables(p, box, "cpu")
        for k in ("sx", "sy", "sz"):
            # libdevice sin vs libm sin: both are within 1-2 ulp of the true value
            torch.testing.assert_close(got[k].cpu(), want[k], rtol=0, atol=4e-16)
        for a in "xyz":
            assert torch.equal(got["rs" + a].cpu(), 1.0 / got["s" + a].cpu().abs())


def test_layer0_matches_table_product():
    p = w.Problem(N=41, Lx=1.0, Ly=2.0, Lz=3.0, T=0.1, timesteps=4)
    dec = Decomposition.create(p.N, 4)
    for r in range(4):
        box, lay = dec.box(r), dec.layout(r)
        tabs = K.tables(p, box, DEV)
        u = K.alloc_field(lay, DEV)
        K.layer0(p, box, lay, tabs, u)
        want = R.analytic(p, {k: v for k, v in tabs.items()}, 0)
        assert torch.equal(K.owned_view(u, box, lay), want)
        # nothing outside the owned cells was written
        chk = u.clone()
        K.owned_view(chk, box, lay).zero_()
        assert float(chk.abs().max()) == 0.0


CASES = [
    # (N, nranks, dims, ranks to test)
    (23, 1, None, [0]),
    (23, 8, None, [0, 3, 5, 7]),
    (26, 27, (3, 3, 3), list(range(27))),  # every boundary / neighbour combination
    (70, 2, (1, 1, 2), [0, 1]),            # more than one z tile
    (45, 4, (1, 4, 1), [1, 2]),
]


@pytest.mark.parametrize("N,nranks,dims,ranks", CASES)
@pytest.mark.parametrize("lengths", [(1.0, 1.0, 1.0), (1.0, 2.0, _lib.PI_FULL)])
def test_step_matches_cpu_twin_and_torch_reference(N, nranks, dims, ranks, lengths):
    p = w.Problem(N=N, Lx=lengths[0], Ly=lengths[1], Lz=lengths[2], T=0.1, timesteps=4)
    dec = Decomposition.create(p.N, nranks, dims)
    g = torch.Generator().manual_seed(7)
    for r in ranks:
        box, lay = dec.box(r), dec.layout(r)
        u1 = torch.rand(lay.elems, dtype=torch.float64, generator=g)
        u0 = torch.rand(lay.elems, dtype=torch.float64, generator=g)
        tabs_d = K.tables(p, box, DEV)
        tabs_h = {k: v.cpu() for k, v in tabs_d.items()}
        for first in (True, False):
            n = 1 if first else 3
            got = u0.to(DEV)
            ga, gr = K.step(p, box, lay, tabs_d, u1.to(DEV), got, n, first)
            torch.cuda.synchronize()
            # (a) plain PyTorch reference on the device tables: a few ulps (no fma in torch)
            want, wa, wr = R.step(p, box, K.owned_view(u1, box, lay, 1).clone(), K.owned_view(u0, box, lay).clone(),
                                  tabs_h, n, first)
            torch.testing.assert_close(K.owned_view(got, box, lay).cpu(), want, rtol=1e-14, atol=1e-14)
            assert abs(ga - wa) <= 1e-12 * max(wa, 1e-300)
            assert abs(gr - wr) <= 1e-9 * max(wr, 1e-300)
            # (b) C++ CPU twin (explicit fma, IEEE divisions): the field must agree bit for bit
            cpu = u0.clone()
            K.step(p, box, lay, None, u1, cpu, n, first)
            assert torch.equal(got.cpu(), cpu), f"rank {r} first={first}: max ulp diff {_ulps(got.cpu(), cpu)}"


@pytest.mark.parametrize("div", [0, 1, 2, 3])
def test_division_modes(div, monkeypatch):
    # mode 0 (multiply by 1/h^2) is only exact for power-of-two h^2; 2 and 3 must equal IEEE division
    lengths = (1.0, 1.0, 1.0) if div == 0 else (1.0, 2.0, _lib.PI_FULL)
    N = 32 if div == 0 else 29
    monkeypatch.setenv("WAVE3D_DIV", str(div))
    p = w.Problem(N=N, Lx=lengths[0], Ly=lengths[1], Lz=lengths[2], T=0.1, timesteps=4)
    dec = Decomposition.create(p.N, 1)
    box, lay = dec.box(0), dec.layout(0)
    g = torch.Generator().manual_seed(3)
    u1 = torch.rand(lay.elems, dtype=torch.float64, generator=g)
    u0 = torch.rand(lay.elems, dtype=torch.float64, generator=g)
    tabs = K.tables(p, box, DEV)
    got = u0.to(DEV)
    K.step(p, box, lay, tabs, u1.to(DEV), got, 2, False)
    cpu = u0.clone()
    K.step(p, box, lay, None, u1, cpu, 2, False)
    if div == 1:
        assert _ulps(got.cpu(), cpu) <= 1  # faithful, not proven correctly rounded
    else:
        assert torch.equal(got.cpu(), cpu)


def test_zero_walls_and_faces():
    p = w.Problem(N=19, T=0.1, timesteps=2)
    dec = Decomposition.create(p.N, 8)
    for r in range(8):
        box, lay = dec.box(r), dec.layout(r)
        h = torch.rand(lay.elems, dtype=torch.float64)
        d = h.to(DEV)
        K.zero_walls(box, lay, d)
        K.zero_walls(box, lay, h)
        assert torch.equal(d.cpu(), h)
        for f in range(6):
            if box.nbr[f] < 0:
                continue
            assert torch.equal(K.pack_face(box, lay, d, f).cpu(), K.pack_face(box, lay, h, f))
            buf = torch.rand(K.face_elems(box)[f], dtype=torch.float64)
            K.unpack_face(box, lay, d, f, buf.to(DEV))
            K.unpack_face(box, lay, h, f, buf)
            assert torch.equal(d.cpu(), h)


def test_field_error_two_stage_reduction():
    p = w.Problem(N=50, Lx=1.0, Ly=2.0, Lz=3.0, T=0.1, timesteps=4)
    dec = Decomposition.create(p.N, 2)
    for r in range(2):
        box, lay = dec.box(r), dec.layout(r)
        tabs = K.tables(p, box, DEV)
        u = torch.rand(lay.elems, dtype=torch.float64)
        d = u.to(DEV)
        scratch = K.field_error_scratch(box, DEV)
        for _ in range(2):  # the scratch is left ready for the next call
            a, rr = K.field_error(p, box, lay, tabs, d, 2, scratch=scratch)
            wa, wr = R.field_error(p, box, K.owned_view(u, box, lay), {k: v.cpu() for k, v in tabs.items()}, 2)
            assert a == wa
            assert abs(rr - wr) <= 1e-12 * wr


def test_fused_rel_error_filter_equals_exact_division():
    """The candidate filter must return exactly max(a/|f|) — compare with the stand-alone reduction
    (which divides every cell) on the layer the fused kernel just produced."""
    p = w.Problem(N=64, T=0.025, timesteps=20)
    dec = Decomposition.create(p.N, 1)
    box, lay = dec.box(0), dec.layout(0)
    tabs = K.tables(p, box, DEV)
    u0, u1 = K.alloc_field(lay, DEV), K.alloc_field(lay, DEV)
    K.layer0(p, box, lay, tabs, u0)
    for f in (0, 1):
        K.unpack_face(box, lay, u0, f ^ 1, K.pack_face(box, lay, u0, f))
    a, r = K.step(p, box, lay, tabs, u0, u1, 1, True)
    fa, fr = K.field_error(p, box, lay, tabs, u1, 1)
    assert (a, r) == (fa, fr)


def test_step_rejects_plans_whose_end_chunks_miss_the_seam_planes():
    """The x-face flags (and the periodic wrap inside one box) are released when the chunks at the two ends of a
    tile column are done, so those chunks have to contain the seam planes send_x[0] / send_x[1]; cuts must
    increase strictly and the ring depths are bounded. A caller-supplied plan that breaks this is refused."""
    p = w.Problem(N=40, T=0.05, timesteps=4)
    dec = Decomposition.create(p.N, 1)
    box, lay = dec.box(0), dec.layout(0)
    tabs = K.tables(p, box, DEV)
    u0, u1 = K.alloc_field(lay, DEV), K.alloc_field(lay, DEV)
    K.layer0(p, box, lay, tabs, u0)
    good = K.plan_for(box)
    K.step(p, box, lay, tabs, u0, u1, 1, first=True, plan=good)  # the planner's own plan is fine

    def broken(edit):
        pl = _lib.Plan.from_buffer_copy(bytes(good))
        edit(pl)
        with pytest.raises(RuntimeError, match="tile plan"):
            K.step(p, box, lay, tabs, u0, u1, 1, first=True, plan=pl)

    def one_plane_end_chunk(pl):
        pl.nchunks = 3
        pl.xcut[0], pl.xcut[1], pl.xcut[2], pl.xcut[3] = 0, 1, 20, box.n[0]   # seam plane 1 is not in chunk 0
        pl.order[0], pl.order[1], pl.order[2] = 0, 2, 1
    broken(one_plane_end_chunk)

    def not_increasing(pl):
        pl.nchunks = 3
        pl.xcut[0], pl.xcut[1], pl.xcut[2], pl.xcut[3] = 0, 20, 20, box.n[0]
        pl.order[0], pl.order[1], pl.order[2] = 0, 2, 1
    broken(not_increasing)

    def ring_too_deep(pl):
        pl.s1 = 17
    broken(ring_too_deep)
