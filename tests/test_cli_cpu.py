"""The wave3d program on the CPU backend: stdout and output-file contract of the reference programs
(ref: README.txt:5-11, openmp_sol.cpp:192-229)."""
import os
import subprocess

from wave3d_b200 import _lib

EXE = os.path.join(_lib.REPO, "bin", "wave3d")


def _run(args, tmp_path, **env):
    return subprocess.run([EXE] + [str(a) for a in args], cwd=tmp_path, capture_output=True, text=True,
                          env={**os.environ, "WAVE3D_BACKEND": "cpu", **env}, timeout=300)


def test_cli_cpu_default_steps_and_file(tmp_path):
    r = _run([24, 2, 1, 1, 1], tmp_path)  # T=1, timesteps=20 by default
    assert r.returncode == 0, r.stderr
    assert r.stdout.splitlines()[0] == "C = 0.190986"
    lines = (tmp_path / "output_N24_Np2.txt").read_text().splitlines()
    assert lines[0].startswith("numerical solution calculated in ") and lines[0].endswith("ms")
    assert lines[1] == "max abs and rel errors on layer 0: 0 0"
    assert len(lines) == 1 + 21


def test_cli_cpu_decomposed_full_precision(tmp_path):
    a = _run([32, 1, 1, 1, 1, 0.025, 20], tmp_path, WAVE3D_FULL_PRECISION="1", WAVE3D_OUTPUT="a.txt")
    b = _run([32, 1, 1, 1, 1, 0.025, 20], tmp_path, WAVE3D_FULL_PRECISION="1", WAVE3D_OUTPUT="b.txt", WAVE3D_DIMS="2x1x3")
    assert a.returncode == 0 and b.returncode == 0
    fa = (tmp_path / "a.txt.full").read_text().splitlines()[1:]
    fb = (tmp_path / "b.txt.full").read_text().splitlines()[1:]
    assert fa == fb  # decomposition-invariant to the last digit
    assert fa[20].split()[-2:] == ["1.1278897771305196e-06", "0.0043628395206666334"]
    assert "grid 2x1x3" in b.stdout


def test_cli_usage_error(tmp_path):
    r = _run([32, 1], tmp_path)
    assert r.returncode == 2 and "usage" in r.stderr


def test_cli_progress_mode_matches_batch_mode(tmp_path):
    a = _run([20, 1, 1, 1, 1, 0.025, 4], tmp_path, WAVE3D_FULL_PRECISION="1", WAVE3D_OUTPUT="a.txt", WAVE3D_RANKS="4")
    b = _run([20, 1, 1, 1, 1, 0.025, 4], tmp_path, WAVE3D_FULL_PRECISION="1", WAVE3D_OUTPUT="b.txt", WAVE3D_RANKS="4",
             WAVE3D_PROGRESS="1")
    assert a.returncode == 0 and b.returncode == 0
    assert [l for l in b.stdout.splitlines() if l.startswith("calculating layer")] == \
        [f"calculating layer {n}" for n in range(1, 5)]  # ref: cuda_sol.cpp:338,385
    assert (tmp_path / "a.txt.full").read_text().splitlines()[1:] == (tmp_path / "b.txt.full").read_text().splitlines()[1:]
