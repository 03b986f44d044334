"""CPU backend = oracle. Checked against BASELINE.md §2.3 (the reference's CUDA sources emulated on a
CPU), against an independent naive solver, against the plain-PyTorch restatement, and for
decomposition invariance."""
import ctypes as C

import pytest
import torch

import wave3d_b200 as w
from wave3d_b200 import _lib
from wave3d_b200.ops import kernels as K, reference as R
from wave3d_b200.parallel.decomp import Decomposition

from conftest import BASELINE_CUDA_EMULATED


def _problem(key):
    N, Lx, Ly, Lz, T, steps = key
    f = lambda v: _lib.PI_FULL if v == "pi" else v
    return w.Problem(N=N, Lx=f(Lx), Ly=f(Ly), Lz=f(Lz), T=T, timesteps=steps)


@pytest.mark.parametrize("key", [k for k in BASELINE_CUDA_EMULATED if k[0] <= 64])
def test_matches_emulated_reference_17_digits(key):
    p = _problem(key)
    with w.WaveSolver(p, nranks=1, backend="cpu") as s:
        res = s.run()
    assert res.max_abs[0] == 0.0 and res.max_rel[0] == 0.0
    for layer, (ea, er) in BASELINE_CUDA_EMULATED[key].items():
        assert res.max_abs[layer] == ea  # same libm, same expression order -> same bits
        if er is not None:
            assert res.max_rel[layer] == er


def test_n128_config1_geometry():
    key = (128, 1.0, 1.0, 1.0, 0.025, 20)
    p = _problem(key)
    with w.WaveSolver(p, nranks=8, backend="cpu") as s:
        res = s.run()
    for layer, (ea, er) in BASELINE_CUDA_EMULATED[key].items():
        assert res.max_abs[layer] == ea
        if er is not None:
            assert res.max_rel[layer] == er


@pytest.mark.parametrize("nranks,dims", [(2, None), (3, None), (4, None), (6, None), (8, None), (8, (8, 1, 1)),
                                         (8, (1, 2, 4)), (4, (1, 1, 4)), (12, (2, 3, 2)), (27, (3, 3, 3))])
def test_decomposition_invariance_bit_identical(nranks, dims):
    p = w.Problem(N=37, Lx=1.0, Ly=2.0, Lz=_lib.PI_FULL, T=0.05, timesteps=8)  # odd N, anisotropic, non-divisible
    with w.WaveSolver(p, nranks=1, backend="cpu") as s:
        ref = s.run()
    with w.WaveSolver(p, nranks=nranks, dims=dims, backend="cpu") as s:
        res = s.run()
    assert res.max_abs == ref.max_abs
    assert res.max_rel == ref.max_rel


def test_naive_solver_cross_check():
    p = w.Problem(N=20, Lx=1.0, Ly=1.5, Lz=2.0, T=0.05, timesteps=6)
    n = p.native()
    L = p.timesteps + 1
    a, r = (C.c_double * L)(), (C.c_double * L)()
    assert _lib.lib().wave3d_cpu_solve_naive(C.byref(n), a, r) == 0
    with w.WaveSolver(p, nranks=4, backend="cpu") as s:
        res = s.run()
    assert list(a) == res.max_abs and list(r) == res.max_rel


def test_second_order_convergence():
    errs = []
    for N in (16, 32, 64):
        p = w.Problem(N=N, T=0.025, timesteps=20)
        with w.WaveSolver(p, nranks=1, backend="cpu") as s:
            errs.append(s.run().max_abs[20])
    assert 3.5 < errs[0] / errs[1] < 4.5 and 3.5 < errs[1] / errs[2] < 4.5  # error ~ h^2


def test_truncated_pi_flavour_matches_cpu_reference_values():
    # the CPU variants (PI = 3.1415926535), N=32: layer 1 abs 2.81989e-09, layer 20 1.12788e-06
    p = w.Problem(N=32, T=0.025, timesteps=20, pi=_lib.PI_TRUNC)
    with w.WaveSolver(p, nranks=2, backend="cpu") as s:
        res = s.run()
    from wave3d_b200.utils.spec import format_double as fd
    assert fd(res.max_abs[1]) == "2.81989e-09" and fd(res.max_abs[10]) == "2.82003e-07"
    assert fd(res.max_abs[20]) == "1.12788e-06"
    # the rel column is a roundoff artefact: the CPU variants evaluate sin(2*PI*x/Lx) with a
    # different expression tree, so only ~3 digits carry over
    assert abs(res.max_rel[20] - 1.17328e-06) < 2e-3 * 1.17328e-06


def _setup(p, nranks, dims, rank):
    dec = Decomposition.create(p.N, nranks, dims)
    return dec, dec.box(rank), dec.layout(rank)


@pytest.mark.parametrize("nranks,dims,rank", [(1, None, 0), (8, None, 0), (8, None, 7), (27, (3, 3, 3), 13)])
def test_cpu_step_vs_torch_reference(nranks, dims, rank):
    p = w.Problem(N=23, Lx=1.0, Ly=2.0, Lz=3.0, T=0.1, timesteps=4)
    dec, box, lay = _setup(p, nranks, dims, rank)
    g = torch.Generator().manual_seed(1)
    u1 = torch.rand(lay.elems, dtype=torch.float64, generator=g)
    u0 = torch.rand(lay.elems, dtype=torch.float64, generator=g)
    tabs = R.tables(p, box)
    for first in (True, False):
        U1 = K.owned_view(u1, box, lay, ghosts=1).clone()
        U0 = K.owned_view(u0, box, lay).clone()
        want, wa, wr = R.step(p, box, U1, U0, tabs, 3, first)
        got = u0.clone()
        ga, gr = K.step(p, box, lay, None, u1, got, 3, first)
        G = K.owned_view(got, box, lay)
        torch.testing.assert_close(G, want, rtol=1e-14, atol=1e-14)
        assert abs(ga - wa) <= 1e-13 * max(1.0, wa) and abs(gr - wr) <= 1e-10 * max(1.0, wr)
        # nothing outside the owned update range was touched
        mask = torch.ones_like(got, dtype=torch.bool)
        sl = tuple(slice(box.lo[a], box.hi[a] + 1) for a in range(3))
        K.owned_view(mask, box, lay)[sl] = False
        assert torch.equal(got[mask], u0[mask])


def test_pack_unpack_roundtrip_and_seam_planes():
    p = w.Problem(N=15, T=0.1, timesteps=2)
    dec = Decomposition.create(p.N, 8)
    fields = []
    for r in range(8):
        lay = dec.layout(r)
        u = torch.arange(lay.elems, dtype=torch.float64) + 1e6 * r
        fields.append(u)
    for r in range(8):
        box, lay = dec.box(r), dec.layout(r)
        for f in range(6):
            q = box.nbr[f]
            if q < 0:
                continue
            buf = K.pack_face(box, lay, fields[r], f)
            qb, ql = dec.box(q), dec.layout(q)
            K.unpack_face(qb, ql, fields[q], f ^ 1, buf)
            V = K.owned_view(fields[r], box, lay)
            G = K.owned_view(fields[q], qb, ql, ghosts=1)
            X, Y, Z = box.n
            if f == 0:
                assert torch.equal(G[-1, 1:-1, 1:-1], V[box.send_x[0]])
            elif f == 1:
                assert torch.equal(G[0, 1:-1, 1:-1], V[box.send_x[1]])
            elif f == 3:
                assert torch.equal(G[1:-1, 0, 1:-1], V[:, Y - 1, :])
            elif f == 5:
                assert torch.equal(G[1:-1, 1:-1, 0], V[:, :, Z - 1])


def test_field_error_matches_step_error():
    p = w.Problem(N=18, T=0.05, timesteps=3)
    dec, box, lay = _setup(p, 1, None, 0)
    u0 = K.alloc_field(lay)
    u1 = K.alloc_field(lay)
    K.layer0(p, box, lay, None, u0)
    # periodic self-wrap by hand
    for f in (0, 1):
        K.unpack_face(box, lay, u0, f ^ 1, K.pack_face(box, lay, u0, f))
    a, r = K.step(p, box, lay, None, u0, u1, 1, True)
    fa, fr = K.field_error(p, box, lay, None, u1, 1)
    assert (a, r) == (fa, fr)
    ta, tr = R.field_error(p, box, K.owned_view(u1, box, lay), R.tables(p, box), 1)
    assert abs(ta - a) <= 1e-12 * a


def test_torch_reference_solve_agrees():
    p = w.Problem(N=16, T=0.025, timesteps=5)
    ea, er = R.solve(p)
    with w.WaveSolver(p, backend="cpu") as s:
        res = s.run()
    for n in range(6):
        assert abs(ea[n] - res.max_abs[n]) <= 1e-9 * max(res.max_abs[n], 1e-30) + 1e-18


def test_per_step_driver_matches_run():
    p = w.Problem(N=20, T=0.025, timesteps=5)
    with w.WaveSolver(p, nranks=4, backend="cpu") as s:
        a = s.run()
        b = s.run_per_step()
    assert a.max_abs == b.max_abs and a.max_rel == b.max_rel


def test_post_pass_field_error_equals_fused_error():
    p = w.Problem(N=22, Lx=1.0, Ly=2.0, Lz=3.0, T=0.05, timesteps=6)
    with w.WaveSolver(p, nranks=4, backend="cpu") as s:
        res = s.run()
        assert s.field_error(6) == (res.max_abs[6], res.max_rel[6])
        assert s.field_error(5) == (res.max_abs[5], res.max_rel[5])
