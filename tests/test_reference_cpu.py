"""The reference's own CPU programs, built unmodified against the MPI stand-in, as cross-checks of the
CPU backend — and a functional test of the stand-in itself. Skipped when
baseline/_ref has not been built (it needs /root/reference)."""
import os
import sys

import pytest

import wave3d_b200 as w
from wave3d_b200 import _lib
from wave3d_b200.utils.spec import format_double as fd

sys.path.insert(0, os.path.join(_lib.REPO, "baseline"))
import run_reference as rr  # noqa: E402

needs_ref = pytest.mark.skipif(not rr.available("mpi_new"), reason="baseline/_ref not built")


def _ours(N, steps, pi):
    p = w.Problem(N=N, T=0.025, timesteps=steps, pi=pi)
    with w.WaveSolver(p, nranks=2, backend="cpu") as s:
        return s.run()


@needs_ref
@pytest.mark.parametrize("prog,nranks", [("mpi_new", 1), ("mpi_new", 2), ("mpi_new", 8), ("mpi_sol", 4),
                                         ("hybrid_new", 2), ("hybrid_sol", 3)])
def test_reference_mpi_programs_agree_with_cpu_backend(tmp_path, prog, nranks):
    outs = rr.run_all_ranks_local(prog, [32, 2, 1, 1, 1, 0.025, 20], nranks, str(tmp_path), timeout=300)
    assert all(rc == 0 for rc, _, _ in outs), outs
    assert outs[0][1].splitlines()[0] == "C = 0.0063662"
    files = [f for f in os.listdir(tmp_path) if f.startswith("output_N32")]
    assert len(files) == 1
    ms, ea, er, _ = rr.parse_output(str(tmp_path / files[0]))
    assert len(ea) == 21 and ea[0] == 0.0
    ours = _ours(32, 20, _lib.PI_TRUNC)  # the CPU variants use the truncated PI
    for n in (1, 10, 20):
        assert fd(ours.max_abs[n]) == fd(ea[n])
    assert fd(ea[1]) == "2.81989e-09" and fd(ea[20]) == "1.12788e-06"


@needs_ref
def test_reference_openmp_program(tmp_path):
    import subprocess
    r = subprocess.run([os.path.join(rr.BIN, "openmp_sol"), "32", "4", "1", "1", "1", "0.025", "20"], cwd=tmp_path,
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0
    ms, ea, er, _ = rr.parse_output(str(tmp_path / "output_N32_Np4.txt"))  # the name our CLI also writes
    ours = _ours(32, 20, _lib.PI_TRUNC)
    assert fd(ea[20]) == fd(ours.max_abs[20]) == "1.12788e-06"
