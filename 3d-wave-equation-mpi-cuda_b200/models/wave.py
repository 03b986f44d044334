"""WaveSolver — the flagship solve, all ranks driven from this process.

Replaces the reference program's driver (ref: cuda_sol.cpp:314-443 time loop,
:477-575 set-up) with the native solver object of csrc/wave3d_runtime.cu: one rank
per GPU (or several ranks time-sharing a GPU for tests, or host partitions on the
CPU backend), in-kernel halo push over peer memory, per-layer error maxima on the
device, the whole solve captured in one CUDA graph per GPU.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

from .. import _lib
from ..utils.spec import Problem
from ..utils import report as _report


@dataclass
class SolveResult:
    max_abs: List[float]
    max_rel: List[float]
    solve_ms: float
    init_ms: float = 0.0
    launches: int = 0
    nranks: int = 1
    dims: Sequence[int] = (1, 1, 1)
    cells: int = 0
    timesteps: int = 0
    extra: dict = field(default_factory=dict)

    @property
    def gcells_per_s(self) -> float:
        """(N+1)^3 * timesteps / t_solve / 1e9 — the north-star metric."""
        return self.cells * self.timesteps / (self.solve_ms * 1e-3) / 1e9 if self.solve_ms > 0 else float("nan")


class _DevPtr:
    """Zero-copy view of device memory for torch.as_tensor (CUDA array interface v2)."""

    def __init__(self, ptr: int, nelems: int):
        self.__cuda_array_interface__ = {"shape": (nelems,), "typestr": "<f8", "data": (ptr, False), "version": 2}


class WaveSolver:
    def __init__(self, problem: Problem, nranks: int = 1, dims: Optional[Sequence[int]] = None, backend: str = "cuda",
                 transport: str = "peer", devices: Optional[Sequence[int]] = None, use_graph: bool = True,
                 threads: int = 0, per_step_sync: bool = False, first_rank: int = 0, nlocal: Optional[int] = None,
                 spin_timeout_s: float = 10.0, multistep: bool = True):
        self.problem = problem
        self._p = problem.native()
        o = _lib.SolverOpts()
        L = _lib.lib()
        L.wave3d_solver_opts_default(C.byref(o))
        o.backend = _lib.BACKEND_CUDA if backend == "cuda" else _lib.BACKEND_CPU
        o.transport = _lib.TRANSPORT_PEER if transport == "peer" else _lib.TRANSPORT_STAGED
        o.nranks = nranks
        o.first_rank = first_rank
        o.nlocal = nranks if nlocal is None else nlocal
        if dims:
            for a in range(3):
                o.dims[a] = int(dims[a])
        if devices:
            for i in range(8):
                o.devices[i] = int(devices[i % len(devices)])
        o.use_graph = int(use_graph)
        o.threads = threads
        o.per_step_sync = int(per_step_sync)
        o.spin_timeout_s = spin_timeout_s
        o.multistep = int(multistep)
        self.opts = o
        self.backend = backend
        self.transport = transport
        err = C.create_string_buffer(512)
        self._h = L.wave3d_solver_create(C.byref(self._p), C.byref(o), err, len(err))
        if not self._h:
            raise RuntimeError("wave3d_solver_create: " + err.value.decode())
        self.nranks = nranks
        self.nlocal = o.nlocal
        b = self.box(0)
        self.dims = tuple(b.dims)
        self.init_ms = 0.0

    # ------------------------------------------------------------ lifecycle
    def close(self):
        if getattr(self, "_h", None):
            _lib.lib().wave3d_solver_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _chk(self, rc: int, what: str):
        if rc:
            raise RuntimeError(f"{what} failed (rc={rc}): {_lib.last_error()}")

    # ------------------------------------------------------------- geometry
    def box(self, local: int = 0) -> _lib.Box:
        b = _lib.Box()
        self._chk(_lib.lib().wave3d_solver_box(self._h, local, C.byref(b), None, None), "solver_box")
        return b

    def layout(self, local: int = 0) -> _lib.Layout:
        lay = _lib.Layout()
        self._chk(_lib.lib().wave3d_solver_box(self._h, local, None, C.byref(lay), None), "solver_box")
        return lay

    def plan(self, local: int = 0) -> _lib.Plan:
        pl = _lib.Plan()
        self._chk(_lib.lib().wave3d_solver_box(self._h, local, None, None, C.byref(pl)), "solver_box")
        return pl

    def field(self, local: int, which: int):
        """torch view (no copy) of field 0/1 of a local rank; layer n lives in field n & 1."""
        import torch
        ptr = _lib.lib().wave3d_solver_field(self._h, local, which)
        lay = self.layout(local)
        if self.backend == "cuda":
            dev = self.opts.devices[local % 8]
            dev = dev if dev >= 0 else local % max(1, _lib.lib().wave3d_cuda_device_count())
            return torch.as_tensor(_DevPtr(ptr, lay.elems), device=f"cuda:{dev}")
        buf = (C.c_double * lay.elems).from_address(ptr)
        return torch.frombuffer(buf, dtype=torch.float64)

    def stream(self, local: int = 0) -> int:
        return _lib.lib().wave3d_solver_stream(self._h, local) or 0

    # --------------------------------------------------------------- phases
    def reset(self):
        self._chk(_lib.lib().wave3d_solver_reset(self._h), "solver_reset")

    def enqueue_layer0(self):
        self._chk(_lib.lib().wave3d_solver_enqueue_layer0(self._h), "enqueue_layer0")

    def enqueue_step(self, n: int):
        self._chk(_lib.lib().wave3d_solver_enqueue_step(self._h, n), "enqueue_step")

    def enqueue_step_e2e(self, n: int, pinned_desc_ptr: int):
        self._chk(_lib.lib().wave3d_solver_enqueue_step_e2e(self._h, n, pinned_desc_ptr), "enqueue_step_e2e")

    def step_descriptor(self, n: int):
        d = (C.c_double * 2)()
        _lib.lib().wave3d_step_descriptor(C.byref(self._p), n, d)
        return (d[0], d[1])

    def exchange_local(self, n: int):
        self._chk(_lib.lib().wave3d_solver_exchange_local(self._h, n), "exchange_local")

    def pack(self, local: int, n: int, face: int, buf_ptr: int):
        self._chk(_lib.lib().wave3d_solver_pack(self._h, local, n, face, buf_ptr), "solver_pack")

    def unpack(self, local: int, n: int, face: int, buf_ptr: int):
        self._chk(_lib.lib().wave3d_solver_unpack(self._h, local, n, face, buf_ptr), "solver_unpack")

    def sync(self):
        self._chk(_lib.lib().wave3d_solver_sync(self._h), "solver_sync")

    def layer_error(self, n: int):
        """Blocking read-back of layer n's (max abs, max rel) over the local ranks."""
        e = (C.c_double * 2)()
        self._chk(_lib.lib().wave3d_solver_layer_error(self._h, n, e), "layer_error")
        return e[0], e[1]

    def field_error(self, n: int):
        """(max abs, max rel) of the *stored* layer n against the analytic solution, as a separate post-pass
        (ref: calculate_error(), mpi_sol.cpp:340-372). After a solve the last two layers are resident."""
        e = (C.c_double * 2)()
        self._chk(_lib.lib().wave3d_solver_field_error(self._h, n, e), "solver_field_error")
        return e[0], e[1]

    def errors(self):
        L = self.problem.timesteps + 1
        a = (C.c_double * L)()
        r = (C.c_double * L)()
        self._chk(_lib.lib().wave3d_solver_errors(self._h, a, r), "solver_errors")
        return list(a), list(r)

    @property
    def launches(self) -> int:
        return int(_lib.lib().wave3d_solver_kernel_launches(self._h))

    def status(self) -> int:
        return _lib.lib().wave3d_solver_status(self._h)

    @property
    def multistep_active(self) -> bool:
        """True when layers 2..T run in the persistent multi-layer launch (small boxes, one rank per GPU)."""
        return bool(_lib.lib().wave3d_solver_multistep_active(self._h))

    def export(self, local: int = 0) -> _lib.Export:
        e = _lib.Export()
        self._chk(_lib.lib().wave3d_solver_export(self._h, local, C.byref(e)), "solver_export")
        return e

    def connect(self, exports: Sequence[_lib.Export]):
        arr = (_lib.Export * len(exports))(*exports)
        err = C.create_string_buffer(512)
        rc = _lib.lib().wave3d_solver_connect(self._h, arr, err, len(err))
        if rc:
            raise RuntimeError("solver_connect: " + err.value.decode())

    # ---------------------------------------------------------------- solve
    def run(self) -> SolveResult:
        """Whole solve (layer 0 .. timesteps), device-timed like the reference's
        started->finished events (ref: cuda_sol.cpp:382,422-425)."""
        t = _lib.Timings()
        before = self.launches
        rc = _lib.lib().wave3d_solver_run(self._h, C.byref(t))
        self._chk(rc, "solver_run")
        a, r = self.errors()
        return SolveResult(a, r, t.solve_ms, self.init_ms, self.launches - before, self.nranks, self.dims,
                           self.problem.cells, self.problem.timesteps,
                           {"plan": self.plan(0).as_dict() if self.backend == "cuda" else None,
                            "exchange_ms": t.exchange_ms, "loop_ms": t.loop_ms})

    solve = run

    def measure_phases(self) -> SolveResult:
        """Whole solve plus the reference's phase totals (ref: cuda_sol.cpp:438-441): `extra` carries
        exchange_ms (peer path: time CTAs waited for a neighbour rank's flag), error_ms (cost of the fused
        error evaluation, from one extra solve without it), hostdev_ms (0) and loop_ms (the rest)."""
        t = _lib.Timings()
        before = self.launches
        self._chk(_lib.lib().wave3d_solver_measure_phases(self._h, C.byref(t)), "solver_measure_phases")
        a, r = self.errors()
        return SolveResult(a, r, t.solve_ms, self.init_ms, self.launches - before, self.nranks, self.dims,
                           self.problem.cells, self.problem.timesteps,
                           {"plan": self.plan(0).as_dict() if self.backend == "cuda" else None,
                            "exchange_ms": t.exchange_ms, "loop_ms": t.loop_ms, "error_ms": t.error_ms,
                            "hostdev_ms": t.hostdev_ms})

    def make_descriptors(self):
        """Pinned host tensor [timesteps+1, 2] with every layer's step inputs (time argument, coefficient)."""
        import torch
        d = torch.zeros((self.problem.timesteps + 1, 2), dtype=torch.float64).pin_memory()
        for n in range(1, self.problem.timesteps + 1):
            d[n, 0], d[n, 1] = self.step_descriptor(n)
        return d

    def run_streamed(self, desc=None, lag: int = 1) -> SolveResult:
        """End-to-end solve through the per-step path: every layer's inputs are copied from pinned host
        memory, every layer's maxima are read back to the host while the solve runs (``lag`` layers behind
        the device; 0 = wait for each layer before launching the next, like the reference). Wall-clock timed."""
        if desc is None:
            desc = self.make_descriptors()
        L = self.problem.timesteps + 1
        out = (C.c_double * (2 * L))()
        t = _lib.Timings()
        before = self.launches
        rc = _lib.lib().wave3d_solver_run_streamed(self._h, desc.data_ptr(), lag, out, C.byref(t))
        self._chk(rc, "solver_run_streamed")
        return SolveResult([out[2 * n] for n in range(L)], [out[2 * n + 1] for n in range(L)], t.solve_ms, self.init_ms,
                           self.launches - before, self.nranks, self.dims, self.problem.cells, self.problem.timesteps,
                           {"h2d_bytes_per_step": 16 * self.nlocal,
                            "d2h_bytes_per_step": (20 if self.multistep_active else 16) * self.nlocal,
                            "lag": lag, "multi_layer_launch": self.multistep_active})

    def run_per_step(self, progress=None) -> SolveResult:
        """Reference-style driver: after every layer the maxima are read back to the host
        (what cuda_sol does with 'calculating layer n' + calculate_errors every step)."""
        import time
        self.reset()
        staged = self.transport == "staged" or self.backend != "cuda"
        a, r = [0.0], [0.0]
        before = self.launches
        t0 = time.perf_counter()
        self.enqueue_layer0()
        if staged:
            self.exchange_local(0)
        for n in range(1, self.problem.timesteps + 1):
            if progress:
                progress(n)
            self.enqueue_step(n)
            if staged and n < self.problem.timesteps:
                self.exchange_local(n)
            ea, er = self.layer_error(n)
            a.append(ea)
            r.append(er)
        self.sync()
        ms = (time.perf_counter() - t0) * 1e3
        return SolveResult(a, r, ms, self.init_ms, self.launches - before, self.nranks, self.dims, self.problem.cells,
                           self.problem.timesteps)

    # --------------------------------------------------------------- report
    def report(self, res: SolveResult, precision: int = 0) -> str:
        return _report.format_report(self.problem, res.max_abs, res.max_rel, res.solve_ms, res.init_ms,
                                     loop_ms=res.extra.get("loop_ms"), exchange_ms=res.extra.get("exchange_ms", 0.0),
                                     error_ms=res.extra.get("error_ms", 0.0),
                                     precision=precision, cuda_style=self.backend == "cuda")

    def write_report(self, res: SolveResult, path: Optional[str] = None, precision: int = 0) -> str:
        path = path or _report.output_name(self.problem.N, self.problem.Np)
        with open(path, "w") as f:
            f.write(self.report(res, precision))
        return path
