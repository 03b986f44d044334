"""Inter-process plumbing over torch.distributed (NCCL on GPUs, gloo on CPU).

torch.distributed is used for what MPI did outside the hot loop in the reference
(ref: cuda_sol.cpp:448-451 init, :501-508 local rank -> GPU, :575 barrier,
:431-432 reduce of the maxima) and for the *staged* halo transport. On the default
*peer* transport it only carries the cudaIpc handles once: after that the halo
faces and the epoch flags move inside the step kernel over NVLink peer memory.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List

import torch
import torch.distributed as dist

from .. import _lib


def init_process_group_from_env(backend: str | None = None) -> None:
    """Initialise torch.distributed from torchrun's environment (RANK, WORLD_SIZE, MASTER_*)."""
    if dist.is_initialized():
        return
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl" and ranks_share_a_gpu():
        # more processes than GPUs on this node (the reference's MPS launch mode, ref: cuda_sol.cpp:517-519,
        # README.txt:44-46): NCCL refuses two ranks on one device, so the plumbing runs over gloo and the halo
        # faces are staged through the host — which is what the reference does with MPI
        import warnings
        warnings.warn("wave3d: more ranks than GPUs on this node: using gloo (host-staged halo) instead of NCCL")
        backend = "gloo"
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29533")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    kw = {}
    if backend == "nccl":
        dev = local_device()
        torch.cuda.set_device(dev)
        try:
            kw["device_id"] = torch.device("cuda", dev)
        except Exception:
            pass
    dist.init_process_group(backend, rank=rank, world_size=world, **kw)


def local_device() -> int:
    """rank -> GPU round-robin on the node (ref: cudaSetDevice(local_rank % num_devices), cuda_sol.cpp:519)."""
    n = max(1, torch.cuda.device_count())
    return int(os.environ.get("LOCAL_RANK", os.environ.get("RANK", "0"))) % n


def ranks_share_a_gpu() -> bool:
    """True when torchrun started more local ranks than this node has GPUs (rank -> GPU is round-robin)."""
    if not torch.cuda.is_available():
        return False
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")))
    return local_world > torch.cuda.device_count()


def all_gather_exports(e: _lib.Export) -> List[_lib.Export]:
    """All-gather one plain-bytes export blob per rank, in rank order."""
    raw = bytes(e)
    blobs = [None] * dist.get_world_size()
    dist.all_gather_object(blobs, raw)
    return [_lib.Export.from_buffer_copy(b) for b in blobs]


def all_agree(ok: bool, device=None) -> bool:
    t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(t.item())


def max_over_ranks(values, device=None) -> list:
    t = torch.tensor(list(values), dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


class StagedHalo:
    """pack -> isend/irecv -> unpack of the six faces of one rank (ref: exchange(),
    cuda_sol.cpp:230-312, without the host staging: buffers stay on the device and
    travel by NCCL; on CPU they travel by gloo)."""

    def __init__(self, solver, rank: int, device):
        self.s, self.rank, self.device = solver, rank, device
        self.box = solver.box(0)
        from .decomp import Decomposition
        fe = (C.c_int64 * 6)()
        _lib.lib().wave3d_face_elems(C.byref(self.box), fe)
        self.send = [torch.empty(fe[f], dtype=torch.float64, device=device) if self.box.nbr[f] >= 0 else None
                     for f in range(6)]
        self.recv = [torch.empty(fe[f], dtype=torch.float64, device=device) if self.box.nbr[f] >= 0 else None
                     for f in range(6)]
        # device buffers under a CPU process group (gloo: ranks that share a GPU): the faces travel through
        # pinned host mirrors, like the reference's pinned Matrix buffers (ref: cuda_sol.cpp:73-123, 236-304)
        self.host_staged = torch.device(device).type == "cuda" and dist.get_backend() != "nccl"
        if self.host_staged:
            self.hsend = [torch.empty(t.shape, dtype=t.dtype).pin_memory() if t is not None else None for t in self.send]
            self.hrecv = [torch.empty(t.shape, dtype=t.dtype).pin_memory() if t is not None else None for t in self.recv]

    def exchange(self, n: int):
        b, s = self.box, self.s
        for f in range(6):
            if b.nbr[f] >= 0:
                s.pack(0, n, f, self.send[f].data_ptr())
        send, recv = self.send, self.recv
        if self.host_staged:
            for f in range(6):
                if send[f] is not None:
                    self.hsend[f].copy_(send[f], non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
            send, recv = self.hsend, self.hrecv
        ops = []
        # Post in the sender's face order on both sides so that two messages between the
        # same pair (x faces when dims[0] == 2) match up under NCCL's in-order semantics.
        for sf in range(6):
            q = b.nbr[sf]
            if q >= 0 and q != self.rank:
                ops.append(dist.P2POp(dist.isend, send[sf], q, tag=sf))
        for sf in range(6):
            g = sf ^ 1  # my ghost face g receives the neighbour's send face sf
            q = b.nbr[g]
            if q < 0:
                continue
            if q == self.rank:  # periodic self-wrap
                recv[g].copy_(send[sf])
            else:
                ops.append(dist.P2POp(dist.irecv, recv[g], q, tag=sf))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        if self.host_staged:
            for g in range(6):
                if recv[g] is not None:
                    self.recv[g].copy_(recv[g], non_blocking=True)
        for g in range(6):
            if b.nbr[g] >= 0:
                s.unpack(0, n, g, self.recv[g].data_ptr())
