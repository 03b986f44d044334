"""bench.py prints ONE JSON line with the agreed keys (own arm), and the reference arm either reports the
same metric or says why it is unavailable and still exits 0."""
import json
import os
import subprocess
import sys

import pytest

from wave3d_b200 import _lib

BENCH = os.path.join(_lib.REPO, "bench.py")


def _run(args, timeout=600):
    r = subprocess.run([sys.executable, BENCH] + args, capture_output=True, text=True, timeout=timeout, cwd=_lib.REPO)
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    return r, (json.loads(lines[-1]) if lines else None)


def test_reference_arm_never_fails_hard():
    r, d = _run(["--impl", "reference", "--grid", "32", "--steps", "4", "--warmup", "3"])
    assert r.returncode == 0 and d is not None and d["impl"] == "reference"
    assert "unavailable" in d or (d["value"] > 0 and d["unit"] == "Gcell-updates/s")


@pytest.mark.gpu
def test_own_arm_json_contract():
    r, d = _run(["--grid", "96", "--steps", "6", "--warmup", "3", "--repeats", "2"])
    assert r.returncode == 0, r.stderr[-2000:]
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "roofline"):
        assert k in d, k
    assert d["n_gpus"] == 1 and d["steps"] == 6 and d["warmup"] == 3 and d["higher_is_better"] is True
    assert d["unit"] == "Gcell-updates/s" and d["dtype"] == "fp64" and d["gpu_launches"] > 0
    assert abs(d["value"] - 97 ** 3 * 6 / (d["ms_per_step"] * 6 * 1e-3) / 1e9) < 1e-6 * d["value"]
    e = d["e2e"]
    assert e["value"] > 0 and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0
    assert e["value"] <= d["value"] * 1.05  # end to end cannot beat the device-timed solve
    assert e["maxima_equal_device_timed_solve"] is True
    assert {"sm_mhz", "sm_max_mhz", "reasons"} <= set(d["clocks"])
    assert 0 < d["detail"]["max_abs_error_last_layer"] < 1e-6
    rf = d["roofline"]
    assert rf["bound"] == "hbm" and rf["halo_bytes_per_step"] == 0 and 0 < rf["fraction"] < 1.5


def test_both_arms_name_the_problem_identically():
    """`config` must be byte-identical in both arms (the driver compares it); arm-specific text lives in `detail`."""
    sys.path.insert(0, _lib.REPO)
    import bench
    sys.path.insert(0, os.path.join(_lib.REPO, "baseline"))
    import run_reference as rr
    c = bench.problem_config(512, "1", bench.TAU, 20, rr.dims_create(8), 513.0 ** 3)
    assert c["model"] == "wave3d N=512 L=1 tau=0.00125 timesteps=20" and c["parallelism"] == "2x2x2"
    assert c["global_batch"] == 513 ** 3 and c["seq_len"] == 20
    # the reference's MPI_Dims_create grid equals the native planner's for every rank count of interest
    from wave3d_b200.parallel.decomp import dims_create
    for n in (1, 2, 3, 4, 6, 8, 12, 16):
        assert rr.dims_create(n) == dims_create(n)
    # halo bytes: N=512 on (2,2,2): one x face pair... rank with the largest faces sends 3 faces of <=257x257 cells
    hb = bench.halo_bytes_per_step(512, (2, 2, 2))
    assert 3 * 256 * 256 * 8 <= hb <= 4 * 257 * 257 * 8
    assert bench.halo_bytes_per_step(512, (1, 1, 1)) == 0
    h2d, d2h = rr.staged_bytes_per_step(512, (1, 1, 1))
    assert h2d == 6 * 515 * 515 * 8 and d2h > h2d
