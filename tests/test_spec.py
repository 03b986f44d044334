"""CLI contract, derived constants, report format (SURVEY §2.6, §4.2 'unit (CPU, no GPU)')."""
import math

import pytest

import wave3d_b200 as w
from wave3d_b200 import _lib
from wave3d_b200.utils import report, spec


def test_parse_full_argv():
    p = w.parse_args(["128", "1", "1", "2", "3", "0.025", "20"])
    assert (p.N, p.Np, p.Lx, p.Ly, p.Lz, p.T, p.timesteps) == (128, 1, 1.0, 2.0, 3.0, 0.025, 20)


def test_parse_defaults_and_pi_literal():
    p = w.parse_args(["64", "4", "pi", "pi", "1"])
    assert p.T == 1.0 and p.timesteps == 20  # ref: cuda_sol.cpp:461-464
    assert p.Lx == _lib.PI_FULL and p.Ly == _lib.PI_FULL and p.Lz == 1.0
    q = w.parse_args(["64", "4", "pi", "1", "1"], pi=_lib.PI_TRUNC)
    assert q.Lx == 3.1415926535  # CPU variants' constant


def test_parse_errors_are_reported_not_crashes():
    with pytest.raises(ValueError):
        w.parse_args(["64", "1", "1"])  # the reference crashes in std::stoi here
    with pytest.raises(ValueError):
        w.parse_args(["64", "1", "foo", "1", "1"])
    with pytest.raises(ValueError):
        w.parse_args(["2", "1", "1", "1", "1"])


def test_derived_constants():
    p = w.Problem(N=128, Lx=1, Ly=2, Lz=3, T=0.025, timesteps=20)
    n = p.native()
    PI = _lib.PI_FULL
    assert n.a2 == 1 / (4 * PI * PI)
    assert n.a_t == 0.5 * math.sqrt(4 / 1 + 1 / 4 + 1 / 9)
    assert n.tau == 0.025 / 20
    assert (n.hx, n.hy, n.hz) == (1 / 128, 2 / 128, 3 / 128)
    assert n.courant == math.sqrt(n.a2) * n.tau / min(n.hx, n.hy, n.hz)


def test_courant_line_matches_reference_format():
    # N=128 L=1 T=0.025 20 steps -> C = 0.0254648 ; §2.2: N=32 -> 0.0063662
    assert spec.courant_line(w.Problem(N=128, T=0.025, timesteps=20)) == "C = 0.0254648"
    assert spec.courant_line(w.Problem(N=32, T=0.025, timesteps=20)) == "C = 0.0063662"
    assert spec.courant_line(w.Problem(N=64, Lx=_lib.PI_FULL, Ly=_lib.PI_FULL, Lz=_lib.PI_FULL, T=0.025,
                                       timesteps=20)) == "C = 0.00405285"


def test_format_double_is_iostream_default():
    assert spec.format_double(2.8201634361124661e-09) == "2.82016e-09"
    assert spec.format_double(0.0) == "0"
    assert spec.format_double(0.0043628395206666334) == "0.00436284"
    assert spec.format_double(float("inf")) == "inf"


def test_output_names():
    assert report.output_name(128, 8) == "output_N128_Np8.txt"  # ref: README.txt:11
    assert report.output_name(128, 8, 4) == "output_N128_Np8_Ng4_cuda.txt"  # ref: cuda_sol.cpp:535


def test_report_lines_cuda_style():
    p = w.Problem(N=32, T=0.025, timesteps=2)
    txt = report.format_report(p, [0.0, 2.8201634361124661e-09, 1.12806e-08], [0.0, 1.0969951822587042e-05, 4.4e-05],
                               solve_ms=12.3456789, init_ms=41.9)
    lines = txt.splitlines()
    assert lines[0] == "initialization done in 41ms"
    assert lines[1] == "numerical solution calculated in 12.3457ms"
    assert lines[2] == "max abs and rel errors on layer 0: 0 0"
    assert lines[3] == "max abs and rel errors on layer 1: 2.82016e-09 1.097e-05"
    assert lines[5].startswith("total host-device exchange time: ")
    assert lines[6].startswith("total loop time: ")
    assert lines[7].startswith("total MPI exchange time: ")
    assert lines[8].startswith("total error calculation time: ")
    assert len(lines) == 9


def test_report_cpu_style_and_full_precision():
    p = w.Problem(N=32, T=0.025, timesteps=1)
    txt = report.format_report(p, [0.0, 2.8201634361124661e-09], [0.0, 1.0969951822587042e-05], solve_ms=154.9,
                               cuda_style=False, precision=17)
    lines = txt.splitlines()
    assert lines[0] == "numerical solution calculated in 154ms"  # ref: openmp_sol.cpp prints an integer
    assert lines[2] == "max abs and rel errors on layer 1: 2.8201634361124661e-09 1.0969951822587042e-05"
    assert len(lines) == 3
