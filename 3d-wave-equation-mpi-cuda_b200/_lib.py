"""ctypes binding of libwave3d.so (include/wave3d.h).

The shared library is built in-tree (``make`` or :func:`build`) and loaded from
``<repo>/lib``; nothing is installed into site-packages so the GPU box sees the
same file the tests exercised.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import threading

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(REPO, "lib", "libwave3d.so")

PI_FULL = 3.14159265358979323846
PI_TRUNC = 3.1415926535

BACKEND_CPU, BACKEND_CUDA = 0, 1
TRANSPORT_PEER, TRANSPORT_STAGED = 0, 1
FACES = ("x-", "x+", "y-", "y+", "z-", "z+")


class Problem(C.Structure):
    _fields_ = [("N", C.c_int32), ("Np", C.c_int32), ("timesteps", C.c_int32), ("reserved", C.c_int32),
                ("Lx", C.c_double), ("Ly", C.c_double), ("Lz", C.c_double), ("T", C.c_double), ("pi", C.c_double),
                ("a2", C.c_double), ("a_t", C.c_double), ("tau", C.c_double),
                ("hx", C.c_double), ("hy", C.c_double), ("hz", C.c_double), ("courant", C.c_double)]


class Box(C.Structure):
    _fields_ = [("rank", C.c_int32), ("nranks", C.c_int32), ("dims", C.c_int32 * 3), ("coords", C.c_int32 * 3),
                ("n", C.c_int32 * 3), ("o", C.c_int32 * 3), ("nbr", C.c_int32 * 6),
                ("lo", C.c_int32 * 3), ("hi", C.c_int32 * 3), ("elo", C.c_int32 * 3), ("ehi", C.c_int32 * 3),
                ("send_x", C.c_int32 * 2), ("N", C.c_int32), ("reserved", C.c_int32)]


class Layout(C.Structure):
    _fields_ = [("pitch", C.c_int64), ("rows", C.c_int64), ("planes", C.c_int64), ("plane_stride", C.c_int64),
                ("elems", C.c_int64), ("oz", C.c_int32), ("reserved", C.c_int32)]


class Timings(C.Structure):
    _fields_ = [("init_ms", C.c_double), ("solve_ms", C.c_double), ("loop_ms", C.c_double),
                ("exchange_ms", C.c_double), ("hostdev_ms", C.c_double), ("error_ms", C.c_double)]


MAX_CHUNKS = 32  # WAVE3D_MAX_CHUNKS


class Plan(C.Structure):
    _fields_ = [("tz", C.c_int32), ("ty", C.c_int32), ("rj", C.c_int32), ("warps", C.c_int32),
                ("nty", C.c_int32), ("ntz", C.c_int32), ("nchunks", C.c_int32), ("s1", C.c_int32), ("s0", C.c_int32),
                ("smem_bytes", C.c_int32), ("ctas", C.c_int32), ("slots", C.c_int32), ("efficiency", C.c_double),
                ("xcut", C.c_int32 * (MAX_CHUNKS + 2)), ("order", C.c_int32 * MAX_CHUNKS)]

    def as_dict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_ if k not in ("xcut", "order")}
        d["xcut"] = list(self.xcut[:self.nchunks + 1])
        d["order"] = list(self.order[:self.nchunks])
        return d


class Peers(C.Structure):
    _fields_ = [("dst", C.c_void_p * 6), ("sa", C.c_int64 * 6), ("sb", C.c_int64 * 6),
                ("flag_out", C.c_void_p * 6), ("flag_in", C.c_void_p * 6), ("root_err", C.c_void_p)]


class SolverOpts(C.Structure):
    _fields_ = [("backend", C.c_int32), ("transport", C.c_int32), ("dims", C.c_int32 * 3), ("nranks", C.c_int32),
                ("first_rank", C.c_int32), ("nlocal", C.c_int32), ("devices", C.c_int32 * 8),
                ("use_graph", C.c_int32), ("threads", C.c_int32), ("per_step_sync", C.c_int32),
                ("verbose", C.c_int32), ("multistep", C.c_int32), ("reserved", C.c_int32),
                ("spin_timeout_s", C.c_double)]


class Export(C.Structure):
    _fields_ = [("ipc_handle", C.c_ubyte * 64), ("off_field", C.c_int64 * 2), ("off_flags", C.c_int64),
                ("off_err", C.c_int64), ("slab_bytes", C.c_int64), ("layout", Layout), ("box", Box),
                ("device", C.c_int32), ("pid", C.c_int32), ("plan_nchunks", C.c_int32),
                ("plan_xcut", C.c_int32 * (MAX_CHUNKS + 2)), ("reserved", C.c_int32)]


SCRATCH_WORDS = 16
_lib = None
_lock = threading.Lock()


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile libwave3d.so and the wave3d CLI for sm_100a with the in-tree Makefile."""
    if force:
        subprocess.run(["make", "-C", REPO, "clean"], check=True, capture_output=not verbose)
    r = subprocess.run(["make", "-C", REPO, "-j8"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("building libwave3d failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    if verbose:
        print(r.stdout[-2000:])
    return LIB_PATH


def _sig(lib):
    P, B, L, D = C.POINTER(Problem), C.POINTER(Box), C.POINTER(Layout), C.POINTER(C.c_double)
    vp, i32, i64, dbl, cs = C.c_void_p, C.c_int, C.c_int64, C.c_double, C.c_char_p
    I3 = C.POINTER(C.c_int32)
    sigs = {
        "wave3d_problem_init": (i32, [P, i32, i32, dbl, dbl, dbl, dbl, i32, dbl, cs, C.c_size_t]),
        "wave3d_parse_args": (i32, [i32, C.POINTER(cs), dbl, P, cs, C.c_size_t]),
        "wave3d_format_courant": (i32, [P, cs, C.c_size_t]),
        "wave3d_format_double": (i32, [dbl, cs, C.c_size_t]),
        "wave3d_dims_create": (None, [i32, I3]),
        "wave3d_parse_dims": (i32, [cs, I3]),
        "wave3d_rank_of": (i32, [I3, I3]),
        "wave3d_coords_of": (None, [I3, i32, I3]),
        "wave3d_box_init": (i32, [i32, I3, i32, B, cs, C.c_size_t]),
        "wave3d_layout_init": (None, [B, L]),
        "wave3d_face_elems": (None, [B, C.POINTER(i64)]),
        "wave3d_format_report": (C.c_size_t, [P, C.POINTER(Timings), D, D, i32, i32, cs, C.c_size_t]),
        "wave3d_output_name": (i32, [i32, i32, i32, cs, C.c_size_t]),
        "wave3d_cpu_tables": (None, [P, B, vp, vp, vp]),
        "wave3d_cpu_coslayer": (dbl, [P, i32]),
        "wave3d_cpu_layer0": (None, [P, B, L, vp, i32]),
        "wave3d_cpu_step": (None, [P, B, L, vp, vp, i32, i32, D, D, i32]),
        "wave3d_cpu_zero_walls": (None, [B, L, vp]),
        "wave3d_cpu_pack_face": (None, [B, L, vp, i32, vp]),
        "wave3d_cpu_unpack_face": (None, [B, L, vp, i32, vp]),
        "wave3d_cpu_field_error": (None, [P, B, L, vp, i32, D, D, i32]),
        "wave3d_cpu_solve": (i32, [P, I3, i32, D, D, C.POINTER(Timings), cs, C.c_size_t]),
        "wave3d_cpu_solve_naive": (i32, [P, D, D]),
        "wave3d_cuda_available": (i32, []),
        "wave3d_cuda_device_count": (i32, []),
        "wave3d_cuda_last_error": (cs, []),
        "wave3d_plan_make": (None, [B, i32, C.POINTER(Plan)]),
        "wave3d_plan_make_style": (None, [B, i32, i32, C.POINTER(Plan)]),
        "wave3d_cuda_tables": (i32, [P, B, vp, vp, vp, vp, vp, vp, vp]),
        "wave3d_cuda_layer0": (i32, [P, B, L, vp, vp, vp, vp, vp]),
        "wave3d_cuda_step": (i32, [P, B, L, C.POINTER(Plan), C.POINTER(vp), vp, vp, i32, i32, C.POINTER(Peers), vp, vp,
                                   dbl, vp]),
        "wave3d_cuda_zero_walls": (i32, [B, L, vp, vp]),
        "wave3d_cuda_pack_face": (i32, [B, L, vp, i32, vp, vp]),
        "wave3d_cuda_unpack_face": (i32, [B, L, vp, i32, vp, vp]),
        "wave3d_cuda_push_faces": (i32, [B, L, vp, C.POINTER(Peers), C.c_uint32, vp]),
        "wave3d_cuda_field_error_scratch_words": (i64, [B]),
        "wave3d_cuda_field_error": (i32, [P, B, L, C.POINTER(vp), vp, i32, vp, vp, vp]),
        "wave3d_cuda_field_error_static": (i32, [P, B, L, C.POINTER(vp), vp, i32, i32, vp, vp]),
        "wave3d_solver_opts_default": (None, [C.POINTER(SolverOpts)]),
        "wave3d_solver_create": (vp, [P, C.POINTER(SolverOpts), cs, C.c_size_t]),
        "wave3d_solver_destroy": (None, [vp]),
        "wave3d_solver_export": (i32, [vp, i32, C.POINTER(Export)]),
        "wave3d_solver_connect": (i32, [vp, C.POINTER(Export), cs, C.c_size_t]),
        "wave3d_solver_box": (i32, [vp, i32, B, L, C.POINTER(Plan)]),
        "wave3d_solver_field": (vp, [vp, i32, i32]),
        "wave3d_solver_stream": (vp, [vp, i32]),
        "wave3d_solver_reset": (i32, [vp]),
        "wave3d_solver_enqueue_layer0": (i32, [vp]),
        "wave3d_solver_enqueue_step": (i32, [vp, i32]),
        "wave3d_step_descriptor": (None, [P, i32, D]),
        "wave3d_solver_enqueue_step_e2e": (i32, [vp, i32, vp]),
        "wave3d_solver_pack": (i32, [vp, i32, i32, i32, vp]),
        "wave3d_solver_unpack": (i32, [vp, i32, i32, i32, vp]),
        "wave3d_solver_exchange_local": (i32, [vp, i32]),
        "wave3d_solver_sync": (i32, [vp]),
        "wave3d_solver_run": (i32, [vp, C.POINTER(Timings)]),
        "wave3d_solver_measure_phases": (i32, [vp, C.POINTER(Timings)]),
        "wave3d_solver_run_streamed": (i32, [vp, vp, i32, D, C.POINTER(Timings)]),
        "wave3d_solver_errors": (i32, [vp, D, D]),
        "wave3d_solver_layer_error": (i32, [vp, i32, D]),
        "wave3d_solver_kernel_launches": (i64, [vp]),
        "wave3d_solver_field_error": (i32, [vp, i32, D]),
        "wave3d_solver_status": (i32, [vp]),
        "wave3d_solver_multistep_active": (i32, [vp]),
        "wave3d_main": (i32, [i32, C.POINTER(cs)]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args


def lib():
    """Load (building first if absent) the native library. Fails loudly: there is no Python fallback."""
    global _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                build()
            _lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
            _sig(_lib)
    return _lib


def last_error() -> str:
    return (lib().wave3d_cuda_last_error() or b"").decode()


def cuda_available() -> bool:
    return bool(lib().wave3d_cuda_available())
