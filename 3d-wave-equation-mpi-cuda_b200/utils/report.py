"""Rank-0 report in the reference's format (ref: cuda_sol.cpp:426-442, 533-536, 572)."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

from .. import _lib
from .spec import Problem


def output_name(N: int, np_: int, ng: int = 0) -> str:
    buf = C.create_string_buffer(256)
    _lib.lib().wave3d_output_name(N, np_, ng, buf, len(buf))
    return buf.value.decode()


def format_report(p: Problem, max_abs: Sequence[float], max_rel: Sequence[float], solve_ms: float = 0.0,
                  init_ms: float = 0.0, loop_ms: Optional[float] = None, exchange_ms: float = 0.0,
                  error_ms: float = 0.0, precision: int = 0, cuda_style: bool = True) -> str:
    n = p.native()
    L = p.timesteps + 1
    assert len(max_abs) == L and len(max_rel) == L
    t = _lib.Timings(init_ms, solve_ms, solve_ms if loop_ms is None else loop_ms, exchange_ms, 0.0, error_ms)
    A = (C.c_double * L)(*max_abs)
    R = (C.c_double * L)(*max_rel)
    need = _lib.lib().wave3d_format_report(C.byref(n), C.byref(t), A, R, precision, int(cuda_style), None, 0)
    buf = C.create_string_buffer(need)
    _lib.lib().wave3d_format_report(C.byref(n), C.byref(t), A, R, precision, int(cuda_style), buf, need)
    return buf.value.decode()


def write_report(path: str, *args, **kw) -> str:
    text = format_report(*args, **kw)
    with open(path, "w") as f:
        f.write(text)
    return text
