"""3-D Cartesian block decomposition: periodic in x, open in y/z, remainder to the
last rank, seam-plane rule (ref: cuda_sol.cpp:477-496, 232).
The arithmetic is native (csrc/wave3d_spec.cpp); the process grid may be forced
(1-D/2-D/3-D) — a capability the reference lacks (BASELINE.json config 4)."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional, Sequence, Tuple

from .. import _lib


def dims_create(nranks: int) -> Tuple[int, int, int]:
    d = (C.c_int32 * 3)()
    _lib.lib().wave3d_dims_create(nranks, d)
    return (d[0], d[1], d[2])


def parse_dims(s: str) -> Tuple[int, int, int]:
    d = (C.c_int32 * 3)()
    if _lib.lib().wave3d_parse_dims(s.encode(), d):
        raise ValueError(f"cannot parse process grid {s!r} (expected e.g. 8x1x1)")
    return (d[0], d[1], d[2])


@dataclass
class Decomposition:
    N: int
    dims: Tuple[int, int, int]

    @staticmethod
    def create(N: int, nranks: int, dims: Optional[Sequence[int]] = None) -> "Decomposition":
        d = tuple(dims) if dims else dims_create(nranks)
        if d[0] * d[1] * d[2] != nranks:
            raise ValueError(f"process grid {d} does not hold {nranks} ranks")
        return Decomposition(N, d)  # type: ignore[arg-type]

    @property
    def nranks(self) -> int:
        return self.dims[0] * self.dims[1] * self.dims[2]

    def box(self, rank: int) -> _lib.Box:
        b = _lib.Box()
        d = (C.c_int32 * 3)(*self.dims)
        err = C.create_string_buffer(256)
        if _lib.lib().wave3d_box_init(self.N, d, rank, C.byref(b), err, len(err)):
            raise ValueError(err.value.decode())
        return b

    def layout(self, rank: int) -> _lib.Layout:
        b = self.box(rank)
        lay = _lib.Layout()
        _lib.lib().wave3d_layout_init(C.byref(b), C.byref(lay))
        return lay

    def rank_of(self, coords: Sequence[int]) -> int:
        d = (C.c_int32 * 3)(*self.dims)
        c = (C.c_int32 * 3)(*coords)
        return _lib.lib().wave3d_rank_of(d, c)

    def coords_of(self, rank: int) -> Tuple[int, int, int]:
        d = (C.c_int32 * 3)(*self.dims)
        c = (C.c_int32 * 3)()
        _lib.lib().wave3d_coords_of(d, rank, c)
        return (c[0], c[1], c[2])

    def face_elems(self, rank: int):
        b = self.box(rank)
        e = (C.c_int64 * 6)()
        _lib.lib().wave3d_face_elems(C.byref(b), e)
        return list(e)
