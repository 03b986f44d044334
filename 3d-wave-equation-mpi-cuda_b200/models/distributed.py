"""DistributedWaveSolver — one rank per process / GPU under torch.distributed.

This is the multi-GPU product path (launched with torchrun, one process per GPU):
  * peer transport (default): each rank exports its slab with cudaIpc, the handles are
    all-gathered once, and from then on the fused step kernel stores its face cells
    directly into the neighbours' ghost cells over NVLink and hand-shakes with epoch
    flags inside the kernel — no NCCL/MPI call per step
    (replaces ref exchange(), cuda_sol.cpp:230-312, and the MPI_Reduce at :431-432).
  * staged transport (fallback and A/B baseline): pack kernels -> NCCL/gloo
    isend/irecv -> unpack kernels every step.
Time is measured on the device and reduced with MAX over ranks.
"""
from __future__ import annotations

import os
import warnings
from typing import Optional, Sequence

import torch
import torch.distributed as dist

from .. import _lib
from ..parallel import dist as pdist
from ..utils.spec import Problem
from .wave import SolveResult, WaveSolver


class DistributedWaveSolver:
    def __init__(self, problem: Problem, dims: Optional[Sequence[int]] = None, backend: Optional[str] = None,
                 transport: str = "auto", use_graph: bool = True, threads: int = 0, spin_timeout_s: float = 10.0,
                 multistep: bool = True):
        if not dist.is_initialized():
            pdist.init_process_group_from_env()
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.problem = problem
        if backend is None:
            backend = "cuda" if torch.cuda.is_available() and _lib.cuda_available() else "cpu"
        self.backend = backend
        self.dev_index = pdist.local_device() if backend == "cuda" else -1
        self.device = torch.device("cuda", self.dev_index) if backend == "cuda" else torch.device("cpu")
        if backend == "cuda":
            torch.cuda.set_device(self.dev_index)
        # collectives run on the process group's native device
        self.coll_device = self.device if dist.get_backend() == "nccl" else torch.device("cpu")
        want_peer = transport in ("auto", "peer") and backend == "cuda"
        # Ranks that share a GPU (more processes than devices, the reference's MPS launch mode, ref:
        # cuda_sol.cpp:517-519): the fused kernels of two processes would spin on each other's flags while
        # time-slicing one device, which can only progress through context switches (and can trip the GPU's
        # watchdog). Such ranks exchange their faces with the staged transport instead — no kernel ever waits.
        self.shared_gpu = False
        if backend == "cuda":
            import socket
            pairs = [None] * self.world
            dist.all_gather_object(pairs, (socket.gethostname(), self.dev_index))
            self.shared_gpu = len(set(pairs)) < len(pairs)
        if self.shared_gpu and want_peer:
            if transport == "peer":
                raise RuntimeError(
                    f"wave3d: {self.world} ranks on {len(set(pairs))} GPU(s): the peer transport needs one GPU per "
                    "process (its kernels wait for each other inside the GPU). Use transport='staged' (or 'auto'), "
                    "or drive several ranks from ONE process: WaveSolver(problem, nranks=..., devices=[...])")
            warnings.warn("wave3d: several ranks share a GPU: using the staged halo transport (no in-kernel waits)")
            want_peer = False
        self.transport = "peer" if want_peer else "staged"
        mk = lambda tr: WaveSolver(problem, nranks=self.world, dims=dims, backend=backend, transport=tr,
                                   devices=[self.dev_index] if backend == "cuda" else None, use_graph=use_graph,
                                   threads=threads, first_rank=self.rank, nlocal=1, spin_timeout_s=spin_timeout_s,
                                   multistep=multistep)
        self.inner = mk(self.transport)
        if want_peer:
            ok, why = True, ""
            try:
                exports = pdist.all_gather_exports(self.inner.export(0))
                self.inner.connect(exports)
            except Exception as e:  # e.g. cudaIpc not permitted in this container
                ok, why = False, str(e)
            if not pdist.all_agree(ok, self.coll_device):
                if transport == "peer":
                    raise RuntimeError(f"peer transport unavailable on rank {self.rank}: {why or 'another rank failed'}")
                warnings.warn(f"wave3d: peer transport unavailable ({why or 'another rank failed'}); using staged NCCL halo")
                self.inner.close()
                self.transport = "staged"
                self.inner = mk("staged")
        self.dims = self.inner.dims
        self.halo = pdist.StagedHalo(self.inner, self.rank, self.device) if self.transport == "staged" else None
        self._ext = torch.cuda.ExternalStream(self.inner.stream(0), device=self.device) if backend == "cuda" else None

    def close(self):
        if getattr(self, "inner", None) is not None:
            if self.backend == "cuda":
                torch.cuda.synchronize(self.device)
            dist.barrier()
            self.halo = None  # pinned / device staging buffers go before the solver and the CUDA context
            self.inner.close()
            self.inner = None

    def _barrier(self):
        if self.backend == "cuda":
            self.inner.sync()
            torch.cuda.synchronize(self.device)
        dist.barrier()

    def _reset(self):
        # late flags of the previous solve must have landed before anyone clears, and
        # everyone must have cleared before anyone publishes again
        self._barrier()
        self.inner.reset()
        self._barrier()

    def run(self) -> SolveResult:
        p = self.problem
        self._reset()
        before = self.inner.launches
        if self.transport == "peer":
            res = self.inner.run()
            ms = res.solve_ms
        else:
            ms = self._run_staged()
        self._barrier()
        a, r = self.inner.errors()
        root_a, root_r = list(a), list(r)
        L = p.timesteps + 1
        red = pdist.max_over_ranks(a + r + [ms], self.coll_device)
        res = SolveResult(red[:L], red[L:2 * L], red[2 * L], 0.0, self.inner.launches - before, self.world, self.dims,
                          p.cells, p.timesteps, {"transport": self.transport, "local_ms": ms,
                                                 "plan": self.inner.plan(0).as_dict() if self.backend == "cuda" else None})
        if self.transport == "peer" and self.rank == 0 and self.world > 1:
            # rank 0's table already holds the global maxima: every rank's last CTA forwarded
            # its maxima with a system-scope atomicMax over peer memory
            res.extra["root_table_matches_allreduce"] = (root_a == res.max_abs and root_r == res.max_rel)
        return res

    solve = run

    def make_descriptors(self):
        return self.inner.make_descriptors()

    def run_streamed(self, desc=None, lag: int = 1) -> SolveResult:
        """End-to-end solve (peer transport): like :meth:`WaveSolver.run_streamed` on every rank — each layer's
        inputs are copied from pinned host memory and its maxima are read back while the solve runs — with the
        per-layer maxima and the wall time reduced with MAX over the ranks."""
        if self.transport != "peer":
            raise RuntimeError("run_streamed needs the peer transport (the staged path is driven step by step)")
        p = self.problem
        self._reset()
        r = self.inner.run_streamed(desc, lag=lag)
        self._barrier()
        L = p.timesteps + 1
        red = pdist.max_over_ranks(list(r.max_abs) + list(r.max_rel) + [r.solve_ms], self.coll_device)
        return SolveResult(red[:L], red[L:2 * L], red[2 * L], 0.0, r.launches, self.world, self.dims, p.cells,
                           p.timesteps, dict(r.extra, local_ms=r.solve_ms))

    def _run_staged(self) -> float:
        p, s = self.problem, self.inner
        if self.backend == "cuda":
            ctx = torch.cuda.stream(self._ext)
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with ctx:
                ev0.record()
                s.enqueue_layer0()
                self.halo.exchange(0)
                for n in range(1, p.timesteps + 1):
                    s.enqueue_step(n)
                    if n < p.timesteps:
                        self.halo.exchange(n)
                ev1.record()
            ev1.synchronize()
            return ev0.elapsed_time(ev1)
        import time
        t0 = time.perf_counter()
        s.enqueue_layer0()
        self.halo.exchange(0)
        for n in range(1, p.timesteps + 1):
            s.enqueue_step(n)
            if n < p.timesteps:
                self.halo.exchange(n)
        return (time.perf_counter() - t0) * 1e3

    def write_report(self, res: SolveResult, path: Optional[str] = None, precision: int = 0) -> Optional[str]:
        """Rank 0 writes output_N{N}_Np{world}.txt like the reference's rank 0 (ref: cuda_sol.cpp:533-536)."""
        if self.rank != 0:
            return None
        from ..utils import report
        path = path or report.output_name(p := self.problem.N, self.world)
        with open(path, "w") as f:
            f.write(report.format_report(self.problem, res.max_abs, res.max_rel, res.solve_ms,
                                         precision=precision, cuda_style=self.backend == "cuda"))
        return path
