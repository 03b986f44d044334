"""End-to-end solves on the GPU: accuracy vs BASELINE.md, decomposition invariance with several
ranks time-sharing one GPU (peer transport: in-kernel ghost stores + epoch flags), staged transport,
CUDA graph vs eager, per-step mode, the CLI binary (SURVEY §4.2 integration / invariance rows)."""
import os
import subprocess

import pytest
import torch

import wave3d_b200 as w
from wave3d_b200 import _lib

from conftest import BASELINE_CUDA_EMULATED

pytestmark = pytest.mark.gpu


def _problem(key):
    N, Lx, Ly, Lz, T, steps = key
    f = lambda v: _lib.PI_FULL if v == "pi" else v
    return w.Problem(N=N, Lx=f(Lx), Ly=f(Ly), Lz=f(Lz), T=T, timesteps=steps)


@pytest.mark.parametrize("key", list(BASELINE_CUDA_EMULATED))
def test_accuracy_vs_emulated_reference(key):
    p = _problem(key)
    with w.WaveSolver(p, nranks=1, backend="cuda") as s:
        res = s.run()
        assert s.status() == 0
    assert res.max_abs[0] == 0.0 and res.max_rel[0] == 0.0
    for layer, (ea, er) in BASELINE_CUDA_EMULATED[key].items():
        # libdevice vs glibc sin/cos differ in the last bits of f: ~1e-7 relative on the abs maximum
        assert abs(res.max_abs[layer] - ea) <= 2e-6 * ea, (layer, res.max_abs[layer], ea)
    # second-order accuracy sanity: abs error grows ~ n^2
    assert 300 < res.max_abs[20] / res.max_abs[1] < 500


def test_gpu_equals_cpu_backend_when_tables_agree():
    """With identical sine tables the GPU field is bit-identical to the CPU twin, so the maxima may only
    differ through libdevice-vs-libm table bits; they must agree to ~1e-7 and be identical in count."""
    p = w.Problem(N=48, Lx=1.0, Ly=2.0, Lz=3.0, T=0.05, timesteps=10)
    with w.WaveSolver(p, backend="cpu") as c, w.WaveSolver(p, backend="cuda") as g:
        rc, rg = c.run(), g.run()
    for n in range(1, 11):
        assert abs(rc.max_abs[n] - rg.max_abs[n]) <= 2e-6 * rc.max_abs[n]


SHAPES = [(2, None), (4, None), (8, None), (8, (8, 1, 1)), (8, (1, 2, 4)), (6, None), (3, (1, 1, 3)), (12, (2, 3, 2))]


@pytest.mark.parametrize("nranks,dims", SHAPES)
def test_decomposition_invariance_peer_transport(nranks, dims):
    """Ranks time-share GPU 0 (launched layer by layer, so no kernel ever waits on a later launch):
    exercises ghost stores, seam planes, walls and the flag protocol. Maxima must be bit-identical."""
    p = w.Problem(N=45, Lx=1.0, Ly=2.0, Lz=_lib.PI_FULL, T=0.05, timesteps=8)
    with w.WaveSolver(p, nranks=1, backend="cuda") as s:
        ref = s.run()
    with w.WaveSolver(p, nranks=nranks, dims=dims, backend="cuda", devices=[0]) as s:
        res = s.run()
        assert s.status() == 0
    assert res.max_abs == ref.max_abs
    assert res.max_rel == ref.max_rel


@pytest.mark.parametrize("nranks,dims", [(2, (1, 1, 2)), (2, (1, 2, 1)), (8, (2, 2, 2))])
@pytest.mark.parametrize("sched", ["0", "8"])
def test_full_tiles_on_rank_faces_keep_the_maxima(nranks, dims, sched, monkeypatch):
    """N=128 gives every rank 64-column tiles that are full (but for, at most, the 64th column: a wall or a
    neighbour's ghost column) AND lie on a y/z rank face: they take the march variant that is the
    predicate-free one plus the ghost stores (WAVE3D_SCHED=8 sends them through the general variant instead).
    Either way the maxima are those of the undivided box, bit for bit."""
    p = w.Problem(N=128, Lx=1.0, Ly=2.0, Lz=_lib.PI_FULL, T=0.025, timesteps=8)
    with w.WaveSolver(p, nranks=1, backend="cuda") as s:
        ref = s.run()
    monkeypatch.setenv("WAVE3D_SCHED", sched)
    with w.WaveSolver(p, nranks=nranks, dims=dims, backend="cuda", devices=[0]) as s:
        res = s.run()
        assert s.status() == 0
    assert res.max_abs == ref.max_abs
    assert res.max_rel == ref.max_rel


@pytest.mark.parametrize("nranks,dims", [(2, None), (8, None), (4, (1, 4, 1))])
def test_staged_transport_equals_peer(nranks, dims):
    p = w.Problem(N=40, T=0.05, timesteps=6)
    with w.WaveSolver(p, nranks=nranks, dims=dims, backend="cuda", devices=[0]) as s:
        a = s.run()
    with w.WaveSolver(p, nranks=nranks, dims=dims, backend="cuda", devices=[0], transport="staged") as s:
        b = s.run()
    assert a.max_abs == b.max_abs and a.max_rel == b.max_rel


def test_graph_equals_eager_and_repeatable():
    p = w.Problem(N=64, T=0.025, timesteps=10)
    with w.WaveSolver(p, nranks=2, backend="cuda", devices=[0], use_graph=True) as s:
        a1, a2 = s.run(), s.run()
    with w.WaveSolver(p, nranks=2, backend="cuda", devices=[0], use_graph=False) as s:
        b = s.run()
    assert a1.max_abs == a2.max_abs == b.max_abs
    assert a1.max_rel == a2.max_rel == b.max_rel
    assert a1.launches == b.launches > p.timesteps


def test_per_step_mode_and_e2e_descriptor():
    p = w.Problem(N=64, T=0.025, timesteps=6)
    with w.WaveSolver(p, backend="cuda") as s:
        ref = s.run()
        per = s.run_per_step()
        assert per.max_abs == ref.max_abs and per.max_rel == ref.max_rel
        # end-to-end flavour: the step descriptor is uploaded from pinned memory every step
        desc = torch.empty((p.timesteps + 1, 2), dtype=torch.float64).pin_memory()
        for n in range(1, p.timesteps + 1):
            desc[n, 0], desc[n, 1] = s.step_descriptor(n)
        s.reset()
        s.enqueue_layer0()
        got = [(0.0, 0.0)]
        for n in range(1, p.timesteps + 1):
            s.enqueue_step_e2e(n, desc[n].data_ptr())
            got.append(s.layer_error(n))
        assert [g[0] for g in got] == ref.max_abs and [g[1] for g in got] == ref.max_rel


def test_field_views_are_live():
    p = w.Problem(N=24, T=0.025, timesteps=3)
    with w.WaveSolver(p, backend="cuda") as s:
        s.run()
        f1 = s.field(0, 1)  # layer 3 lives in field 1
        from wave3d_b200.ops import kernels as K
        box, lay = s.box(0), s.layout(0)
        tabs = K.tables(p, box, "cuda:0")
        a, r = K.field_error(p, box, lay, tabs, f1, 3)
        ea, er = s.errors()
        assert (a, r) == (ea[3], er[3])


def test_cli_binary_writes_reference_format(tmp_path):
    exe = os.path.join(_lib.REPO, "bin", "wave3d")
    out = subprocess.run([exe, "64", "1", "pi", "pi", "pi", "0.025", "20"], cwd=tmp_path, capture_output=True,
                         text=True, env={**os.environ, "WAVE3D_FULL_PRECISION": "1"})
    assert out.returncode == 0, out.stderr
    assert out.stdout.splitlines()[0] == "C = 0.00405285"
    txt = (tmp_path / "output_N64_Np1.txt").read_text().splitlines()
    assert txt[0].startswith("initialization done in ") and txt[0].endswith("ms")
    assert txt[1].startswith("numerical solution calculated in ")
    assert txt[2] == "max abs and rel errors on layer 0: 0 0"
    assert txt[3].startswith("max abs and rel errors on layer 1: 7.15")
    assert len(txt) == 2 + 21 + 4
    full = (tmp_path / "output_N64_Np1.txt.full").read_text().splitlines()
    v = float(full[22].split()[-2])
    assert abs(v - 2.8600414148094444e-08) < 2e-6 * v


def test_cli_multi_rank_on_one_gpu(tmp_path):
    exe = os.path.join(_lib.REPO, "bin", "wave3d")
    out = subprocess.run([exe, "40", "4", "1", "1", "1", "0.025", "5"], cwd=tmp_path, capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert (tmp_path / "output_N40_Np4.txt").exists()


@pytest.mark.parametrize("N,lengths", [(64, (1.0, 1.0, 1.0)), (45, (1.0, 2.0, _lib.PI_FULL)), (130, (1.0, 1.0, 1.0))])
def test_multistep_dataflow_launch_equals_per_step_launches(N, lengths, monkeypatch):
    """Layers 2..T in one persistent launch (per-item epochs instead of kernel boundaries) must give the
    same bits as one launch per layer."""
    p = w.Problem(N=N, Lx=lengths[0], Ly=lengths[1], Lz=lengths[2], T=0.05, timesteps=9)
    monkeypatch.setenv("WAVE3D_MULTISTEP", "auto")
    with w.WaveSolver(p, backend="cuda", multistep=False) as s:
        a = s.run()
    with w.WaveSolver(p, backend="cuda", multistep=True) as s:
        b1 = s.run()
        b2 = s.run()  # epochs / counters are reset between solves
        assert s.status() == 0
        assert b1.launches < a.launches
    assert a.max_abs == b1.max_abs == b2.max_abs
    assert a.max_rel == b1.max_rel == b2.max_rel


def test_staged_transport_reports_exchange_time():
    p = w.Problem(N=48, T=0.05, timesteps=5)
    with w.WaveSolver(p, nranks=2, backend="cuda", devices=[0], transport="staged") as s:
        res = s.run()
        txt = s.report(res)
    assert res.extra["exchange_ms"] > 0.0 and res.extra["loop_ms"] < res.solve_ms
    line = [l for l in txt.splitlines() if l.startswith("total MPI exchange time:")][0]
    assert float(line.split()[-2]) > 0.0


def test_report_phase_totals_on_the_fused_peer_path(tmp_path):
    """The reference prints four phase totals (ref: cuda_sol.cpp:438-441). On the fused peer path they are
    attributed by measurement: exchange = time CTAs waited for a neighbour rank's flag, error = cost of the
    fused error evaluation (one extra solve without it), host-device = 0 by construction."""
    p = w.Problem(N=160, T=0.0125, timesteps=10)
    with w.WaveSolver(p, nranks=2, backend="cuda", devices=[0], transport="peer") as s:
        plain = s.run()
        res = s.measure_phases()
        txt = s.report(res)
    assert res.max_abs == plain.max_abs and res.max_rel == plain.max_rel  # the timing aid leaves the results alone
    assert res.extra["exchange_ms"] > 0.0 and res.extra["error_ms"] > 0.0 and res.extra["hostdev_ms"] == 0.0
    assert 0.0 < res.extra["loop_ms"] < res.solve_ms
    lines = {l.split(":")[0]: float(l.split()[-2]) for l in txt.splitlines() if l.startswith("total ")}
    assert lines["total MPI exchange time"] > 0.0 and lines["total error calculation time"] > 0.0
    assert lines["total host-device exchange time"] == 0.0 and lines["total loop time"] > 0.0
    # the command-line program does the same by default
    exe = os.path.join(_lib.REPO, "bin", "wave3d")
    out = subprocess.run([exe, "160", "2", "1", "1", "1", "0.0125", "10"], cwd=tmp_path, capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert [l for l in out.stdout.splitlines() if l.startswith("calculating layer")] == [
        f"calculating layer {n}" for n in range(1, 11)]  # ref: cuda_sol.cpp:338,385
    cli = {l.split(":")[0]: float(l.split()[-2]) for l in (tmp_path / "output_N160_Np2.txt").read_text().splitlines()
           if l.startswith("total ")}
    assert cli["total error calculation time"] > 0.0 and cli["total MPI exchange time"] > 0.0


@pytest.mark.parametrize("lag", [0, 1, 3])
def test_streamed_solve_equals_batch_solve(lag):
    p = w.Problem(N=56, Lx=1.0, Ly=2.0, Lz=3.0, T=0.05, timesteps=7)
    with w.WaveSolver(p, nranks=2, backend="cuda", devices=[0]) as s:
        ref = s.run()
        got = s.run_streamed(lag=lag)
    assert got.max_abs == ref.max_abs and got.max_rel == ref.max_rel
    assert got.extra["h2d_bytes_per_step"] == 32 and got.solve_ms > 0


def test_streamed_solve_over_the_multi_layer_launch(monkeypatch):
    """Single rank, small box: run_streamed keeps the persistent multi-layer launch; descriptors stream in
    on a copy stream (the kernel waits for each layer's to land), maxima stream out through mapped memory."""
    monkeypatch.setenv("WAVE3D_MULTISTEP", "1")
    p = w.Problem(N=72, Lx=1.0, Ly=2.0, Lz=_lib.PI_FULL, T=0.05, timesteps=11)
    with w.WaveSolver(p, backend="cuda") as s:
        ref = s.run()
        a = s.run_streamed()
        b = s.run_streamed()
        assert s.status() == 0
        assert a.launches < p.timesteps  # one launch for layers 2..T
    assert a.max_abs == ref.max_abs and a.max_rel == ref.max_rel
    assert b.max_abs == ref.max_abs and b.max_rel == ref.max_rel


def test_post_pass_field_error_equals_fused_error():
    p = w.Problem(N=50, Lx=1.0, Ly=2.0, Lz=3.0, T=0.05, timesteps=6)
    for nranks in (1, 4):
        with w.WaveSolver(p, nranks=nranks, backend="cuda", devices=[0]) as s:
            res = s.run()
            assert s.field_error(6) == (res.max_abs[6], res.max_rel[6])
            assert s.field_error(5) == (res.max_abs[5], res.max_rel[5])
            assert s.run().max_abs == res.max_abs  # the scratch borrowed for the post-pass does not disturb a solve


def test_planner_model_ranks_candidate_plans_like_the_gpu_does(monkeypatch):
    """The planner's efficiency model is a handful of calibrated constants; this pins what matters about it:
    for three candidate tile shapes at N=256 its ranking agrees with measured solve times (plans whose
    measured times are within 4 % of each other may swap)."""
    p = w.Problem(N=256, T=0.025, timesteps=20)
    cands = ["14,2", "10,2", "7,1"]
    model, meas = [], []
    for c in cands:
        monkeypatch.setenv("WAVE3D_TILE", c)
        with w.WaveSolver(p, backend="cuda") as s:
            model.append(s.plan(0).efficiency)
            s.run()
            meas.append(min(s.run().solve_ms for _ in range(3)))
    monkeypatch.delenv("WAVE3D_TILE")
    with w.WaveSolver(p, backend="cuda") as s:
        chosen = s.plan(0)
    # the automatic choice is the candidate the model likes best, and it is (one of) the fastest measured
    best_model = max(range(3), key=lambda i: model[i])
    assert f"{chosen.ty},{chosen.rj}" == cands[best_model]
    assert meas[best_model] <= 1.04 * min(meas), (cands, model, meas)
    for i in range(3):
        for j in range(3):
            if model[i] > model[j] and meas[i] > 1.04 * meas[j]:
                raise AssertionError(f"model prefers {cands[i]} over {cands[j]} but it is slower: {model} {meas}")
