"""Problem definition and the reference command-line contract.

``prog N Np Lx Ly Lz [T] [timesteps]`` with the literal ``pi`` accepted for lengths,
defaults ``T=1``, ``timesteps=20`` and the stdout line ``C = ...``
(ref: cuda_sol.cpp:453-475). Parsing and the derived constants live in the native
library (csrc/wave3d_spec.cpp); this is a thin typed wrapper.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Sequence

from .. import _lib


@dataclass(frozen=True)
class Problem:
    N: int
    Np: int = 1
    Lx: float = 1.0
    Ly: float = 1.0
    Lz: float = 1.0
    T: float = 1.0
    timesteps: int = 20
    pi: float = _lib.PI_FULL  # the CUDA variant's constant; PI_TRUNC for the CPU variants

    def native(self) -> _lib.Problem:
        p = _lib.Problem()
        err = C.create_string_buffer(256)
        rc = _lib.lib().wave3d_problem_init(C.byref(p), self.N, self.Np, self.Lx, self.Ly, self.Lz, self.T,
                                            self.timesteps, self.pi, err, len(err))
        if rc:
            raise ValueError(err.value.decode())
        return p

    # derived constants (computed natively so that every consumer sees the same bits)
    @property
    def tau(self) -> float:
        return self.native().tau

    @property
    def courant(self) -> float:
        return self.native().courant

    @property
    def a_t(self) -> float:
        return self.native().a_t

    @property
    def a2(self) -> float:
        return self.native().a2

    @property
    def h(self):
        n = self.native()
        return (n.hx, n.hy, n.hz)

    @property
    def cells(self) -> int:
        return (self.N + 1) ** 3

    @staticmethod
    def from_native(n: _lib.Problem) -> "Problem":
        return Problem(N=n.N, Np=n.Np, Lx=n.Lx, Ly=n.Ly, Lz=n.Lz, T=n.T, timesteps=n.timesteps, pi=n.pi)


def parse_args(argv: Sequence[str], pi: float = _lib.PI_FULL) -> Problem:
    """Parse ``[N, Np, Lx, Ly, Lz, T?, timesteps?]`` (strings, without the program name)."""
    args = [b"wave3d"] + [str(a).encode() for a in argv]
    arr = (C.c_char_p * len(args))(*args)
    p = _lib.Problem()
    err = C.create_string_buffer(512)
    if _lib.lib().wave3d_parse_args(len(args), arr, pi, C.byref(p), err, len(err)):
        raise ValueError(err.value.decode())
    return Problem.from_native(p)


def format_double(v: float) -> str:
    """Default iostream formatting of a double (6 significant digits)."""
    buf = C.create_string_buffer(64)
    _lib.lib().wave3d_format_double(v, buf, len(buf))
    return buf.value.decode()


def courant_line(p: Problem) -> str:
    buf = C.create_string_buffer(64)
    n = p.native()
    _lib.lib().wave3d_format_courant(C.byref(n), buf, len(buf))
    return buf.value.decode()
