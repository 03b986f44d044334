"""Kernel-level entry points of libwave3d on torch tensors.

Every function dispatches on the tensor's device: CUDA tensors run the sm_100a
kernels (csrc/wave3d_kernels.cu), CPU tensors run the OpenMP twins
(csrc/wave3d_cpu.cpp). There is no PyTorch fallback for the CUDA path — if the
native library is missing the import fails.

Reference kernel inventory covered: K1 layer0, K2 layer1, K3
calculate_layer, K4 prepare_layer (walls + seam planes, folded into `step` and
`zero_walls`), K5 copy_to_matrices (`pack_face`), K6 copy_to_grid (`unpack_face`),
K7 calculate_max_errors (`field_error`; fused into `step` on the hot path).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Tuple

import torch

from .. import _lib
from ..utils.spec import Problem

TAB_NAMES = ("sx", "sy", "sz", "rsx", "rsy", "rsz")


def _ptr(t: Optional[torch.Tensor]) -> int:
    return 0 if t is None else t.data_ptr()


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream if t.is_cuda else 0


def _check(rc: int, what: str):
    if rc:
        raise RuntimeError(f"{what} failed: {_lib.last_error()}")


def alloc_field(layout: _lib.Layout, device="cpu") -> torch.Tensor:
    """One zero-initialised time layer in the padded storage layout."""
    return torch.zeros(layout.elems, dtype=torch.float64, device=device)


def storage_view(u: torch.Tensor, layout: _lib.Layout) -> torch.Tensor:
    """[planes, rows, pitch] view of a layer (ghosts and padding included)."""
    return u.view(layout.planes, layout.rows, layout.pitch)


def owned_view(u: torch.Tensor, box: _lib.Box, layout: _lib.Layout, ghosts: int = 0) -> torch.Tensor:
    """[X, Y, Z] view of the owned cells (``ghosts=1``: [X+2, Y+2, Z+2] incl. one ghost layer)."""
    g = ghosts
    v = storage_view(u, layout)
    return v[1 - g:1 + box.n[0] + g, 1 - g:1 + box.n[1] + g, layout.oz - g:layout.oz + box.n[2] + g]


def plan_for(box: _lib.Box, sms: int = 0) -> _lib.Plan:
    p = _lib.Plan()
    _lib.lib().wave3d_plan_make(C.byref(box), sms, C.byref(p))
    return p


def tables(problem: Problem, box: _lib.Box, device="cpu") -> Dict[str, torch.Tensor]:
    """sx, sy, sz (reference argument expressions) and the reciprocals rsx, rsy, rsz."""
    dev = torch.device(device)
    n = problem.native()
    out = {nm: torch.empty(box.n[i % 3], dtype=torch.float64, device=dev) for i, nm in enumerate(TAB_NAMES)}
    if dev.type == "cuda":
        with torch.cuda.device(dev):
            _check(_lib.lib().wave3d_cuda_tables(C.byref(n), C.byref(box), *[_ptr(out[k]) for k in TAB_NAMES],
                                                 _stream(out["sx"])), "wave3d_cuda_tables")
    else:
        _lib.lib().wave3d_cpu_tables(C.byref(n), C.byref(box), _ptr(out["sx"]), _ptr(out["sy"]), _ptr(out["sz"]))
        for a in "xyz":
            out["rs" + a] = 1.0 / out["s" + a].abs()
    return out


def layer0(problem: Problem, box: _lib.Box, layout: _lib.Layout, tabs: Dict[str, torch.Tensor], u0: torch.Tensor):
    n = problem.native()
    if u0.is_cuda:
        with torch.cuda.device(u0.device):
            _check(_lib.lib().wave3d_cuda_layer0(C.byref(n), C.byref(box), C.byref(layout), _ptr(tabs["sx"]),
                                                 _ptr(tabs["sy"]), _ptr(tabs["sz"]), _ptr(u0), _stream(u0)),
                   "wave3d_cuda_layer0")
    else:
        _lib.lib().wave3d_cpu_layer0(C.byref(n), C.byref(box), C.byref(layout), _ptr(u0), 0)
    return u0


def step(problem: Problem, box: _lib.Box, layout: _lib.Layout, tabs: Dict[str, torch.Tensor], u1: torch.Tensor,
         u_io: torch.Tensor, n: int, first: bool = False, plan: Optional[_lib.Plan] = None,
         peers: Optional[_lib.Peers] = None, err: Optional[torch.Tensor] = None,
         scratch: Optional[torch.Tensor] = None) -> Tuple[float, float]:
    """One fused update of layer ``n`` (in place over ``u_io``); returns (max abs, max rel) of this box.

    ``err`` (int64, 2*(timesteps+1)) and ``scratch`` (int32, 16 words, zeroed) may be passed to keep the
    call asynchronous; when omitted they are allocated here and the maxima are read back."""
    pn = problem.native()
    if not u1.is_cuda:
        a, r = C.c_double(0), C.c_double(0)
        _lib.lib().wave3d_cpu_step(C.byref(pn), C.byref(box), C.byref(layout), _ptr(u1), _ptr(u_io), n, int(first),
                                   C.byref(a), C.byref(r), 0)
        return a.value, r.value
    own = err is None
    if err is None:
        err = torch.zeros(2 * (problem.timesteps + 1), dtype=torch.int64, device=u1.device)
    if scratch is None:
        scratch = torch.zeros(_lib.SCRATCH_WORDS, dtype=torch.int32, device=u1.device)
    if plan is None:
        plan = plan_for(box)
    tabarr = (C.c_void_p * 6)(*[_ptr(tabs[k]) for k in TAB_NAMES])
    with torch.cuda.device(u1.device):
        _check(_lib.lib().wave3d_cuda_step(C.byref(pn), C.byref(box), C.byref(layout), C.byref(plan), tabarr, _ptr(u1),
                                           _ptr(u_io), n, int(first), C.byref(peers) if peers is not None else None,
                                           _ptr(err), _ptr(scratch), 5.0, _stream(u1)), "wave3d_cuda_step")
    if own:
        e = err[2 * n:2 * n + 2].view(torch.float64).tolist()
        return e[0], e[1]
    return float("nan"), float("nan")


def zero_walls(box: _lib.Box, layout: _lib.Layout, u: torch.Tensor):
    if u.is_cuda:
        with torch.cuda.device(u.device):
            _check(_lib.lib().wave3d_cuda_zero_walls(C.byref(box), C.byref(layout), _ptr(u), _stream(u)),
                   "wave3d_cuda_zero_walls")
    else:
        _lib.lib().wave3d_cpu_zero_walls(C.byref(box), C.byref(layout), _ptr(u))
    return u


def face_elems(box: _lib.Box):
    e = (C.c_int64 * 6)()
    _lib.lib().wave3d_face_elems(C.byref(box), e)
    return list(e)


def pack_face(box: _lib.Box, layout: _lib.Layout, u: torch.Tensor, face: int,
              buf: Optional[torch.Tensor] = None) -> torch.Tensor:
    if buf is None:
        buf = torch.empty(face_elems(box)[face], dtype=torch.float64, device=u.device)
    if u.is_cuda:
        with torch.cuda.device(u.device):
            _check(_lib.lib().wave3d_cuda_pack_face(C.byref(box), C.byref(layout), _ptr(u), face, _ptr(buf), _stream(u)),
                   "wave3d_cuda_pack_face")
    else:
        _lib.lib().wave3d_cpu_pack_face(C.byref(box), C.byref(layout), _ptr(u), face, _ptr(buf))
    return buf


def unpack_face(box: _lib.Box, layout: _lib.Layout, u: torch.Tensor, face: int, buf: torch.Tensor):
    if u.is_cuda:
        with torch.cuda.device(u.device):
            _check(_lib.lib().wave3d_cuda_unpack_face(C.byref(box), C.byref(layout), _ptr(u), face, _ptr(buf),
                                                      _stream(u)), "wave3d_cuda_unpack_face")
    else:
        _lib.lib().wave3d_cpu_unpack_face(C.byref(box), C.byref(layout), _ptr(u), face, _ptr(buf))
    return u


def field_error_scratch(box: _lib.Box, device) -> torch.Tensor:
    words = _lib.lib().wave3d_cuda_field_error_scratch_words(C.byref(box))
    return torch.zeros(words, dtype=torch.int64, device=device)


def field_error(problem: Problem, box: _lib.Box, layout: _lib.Layout, tabs: Dict[str, torch.Tensor], u: torch.Tensor,
                n: int, scratch: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None):
    """Max abs / rel error of a stored layer against the analytic solution at layer ``n``."""
    pn = problem.native()
    if not u.is_cuda:
        a, r = C.c_double(0), C.c_double(0)
        _lib.lib().wave3d_cpu_field_error(C.byref(pn), C.byref(box), C.byref(layout), _ptr(u), n, C.byref(a),
                                          C.byref(r), 0)
        return a.value, r.value
    if scratch is None:
        scratch = field_error_scratch(box, u.device)
    ret = out is None
    if out is None:
        out = torch.zeros(2, dtype=torch.float64, device=u.device)
    tabarr = (C.c_void_p * 6)(*[_ptr(tabs[k]) for k in TAB_NAMES])
    with torch.cuda.device(u.device):
        _check(_lib.lib().wave3d_cuda_field_error(C.byref(pn), C.byref(box), C.byref(layout), tabarr, _ptr(u), n,
                                                  _ptr(scratch), _ptr(out), _stream(u)), "wave3d_cuda_field_error")
    if ret:
        a, r = out.tolist()
        return a, r
    return out
