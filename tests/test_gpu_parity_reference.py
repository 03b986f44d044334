"""Bit parity with the reference GPU binary (ref: cuda_sol.cpp:431-436 prints the per-layer maxima — the
reference's only oracle). `baseline/_ref/p17/cuda_sol_p17` is the reference built from its own sources with
`out.precision(17)` injected into a COPY (baseline/build_reference.sh), so both programs print %.17g; the
21 "max abs and rel errors on layer n" lines of `bin/wave3d` (WAVE3D_FULL_PRECISION=1 side file) must be
string-identical to the reference's — abs AND rel, every layer, at 1 rank and, when the GPUs are there, at
2 and 8 ranks (the reference ranks share the visible GPUs round-robin, as cuda_sol.cpp:519 does)."""
import os
import subprocess
import sys

import pytest
import torch

from wave3d_b200 import _lib

pytestmark = pytest.mark.gpu
sys.path.insert(0, os.path.join(_lib.REPO, "baseline"))
import run_reference as rr  # noqa: E402

P17 = os.path.join(rr.REF, "p17", "cuda_sol_p17")
CLI = os.path.join(_lib.REPO, "bin", "wave3d")
NGPU = torch.cuda.device_count() if torch.cuda.is_available() else 0


def _error_lines(path):
    return [ln.strip() for ln in open(path) if ln.startswith("max abs and rel errors on layer")]


@pytest.mark.parametrize("N,L,ranks", [(96, "pi", 1), (128, "1", 1), (96, "pi", 2), (128, "1", 2), (96, "pi", 8),
                                       (128, "1", 8)])
def test_per_layer_maxima_are_string_identical_to_the_reference_binary(tmp_path, N, L, ranks):
    if not os.path.exists(P17):
        pytest.skip("baseline/_ref/p17/cuda_sol_p17 is not built (baseline/build_reference.sh needs /root/reference)")
    if NGPU < ranks:
        pytest.skip(f"needs {ranks} GPUs, have {NGPU}")
    K, T = 20, repr(20 * 0.00125)
    ref_dir, our_dir = tmp_path / "ref", tmp_path / "ours"
    ref_dir.mkdir()
    our_dir.mkdir()
    outs = rr.run_all_ranks_local("cuda_sol", [N, 1, L, L, L, T, K], ranks, str(ref_dir), binary=P17,
                                  extra_env={"SHIM_CAP_BYTES": str((N + 3) * (N + 3) * 8 + 4096)})
    assert all(rc == 0 for rc, _, _ in outs), outs
    ref_file = [f for f in os.listdir(ref_dir) if f.startswith("output_N")][0]
    want = _error_lines(ref_dir / ref_file)
    env = dict(os.environ, WAVE3D_FULL_PRECISION="1", WAVE3D_OUTPUT=str(our_dir / "out.txt"), WAVE3D_QUIET="1")
    r = subprocess.run([CLI, str(N), str(ranks), L, L, L, T, str(K)], capture_output=True, text=True, env=env,
                       timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    got = _error_lines(our_dir / "out.txt.full")
    assert len(want) == K + 1 and len(got) == K + 1
    assert got == want
