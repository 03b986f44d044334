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

    def exchange(self, n: int):
        b, s = self.box, self.s
        for f in range(6):
            if b.nbr[f] >= 0:
                s.pack(0, n, f, self.send[f].data_ptr())
        ops = []
        # Post in the sender's face order on both sides so that two messages between the
        # same pair (x faces when dims[0] == 2) match up under NCCL's in-order semantics.
        for sf in range(6):
            q = b.nbr[sf]
            if q >= 0 and q != self.rank:
                ops.append(dist.P2POp(dist.isend, self.send[sf], q, tag=sf))
        for sf in range(6):
            g = sf ^ 1  # my ghost face g receives the neighbour's send face sf
            q = b.nbr[g]
            if q < 0:
                continue
            if q == self.rank:  # periodic self-wrap
                self.recv[g].copy_(self.send[sf])
            else:
                ops.append(dist.P2POp(dist.irecv, self.recv[g], q, tag=sf))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        for g in range(6):
            if b.nbr[g] >= 0:
                s.unpack(0, n, g, self.recv[g].data_ptr())
