"""Allocation-free error evaluation (wave3d_cuda_field_error_static), driven through the C ABI.

Some users of libwave3d cannot allocate device memory at the moment they want an error figure: a
monitor attached to a run that has already reserved all of the GPU's memory, a latency-critical
loop that must not pay the device-wide synchronisation hidden in cudaMalloc/cudaFree, or code that
runs inside another framework's no-allocation region. For them the scratch of the two-stage max
reduction lives in a fixed ~19 KB __device__ array inside the library instead of a caller buffer.

The test calls the new entry point twice in a row (the array still holds the first call's partial
maxima and ticket when the second call starts, and the second call uses a different box and CTA
count) and compares each result with the allocating entry point wave3d_cuda_field_error on the
same data. This file sorts last on purpose: a run that stops at the first failure has then
already executed every other test.
"""
import ctypes as C

import pytest
import torch

import wave3d_b200 as w
from wave3d_b200 import _lib
from wave3d_b200.ops import kernels as K
from wave3d_b200.parallel.decomp import Decomposition

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _static_call(p, box, lay, tabs, u, n, blocks=0):
    """Straight ctypes call of the C entry point; only inputs and the 2-double output are passed."""
    lib = _lib.lib()
    pn = p.native()
    out = torch.zeros(2, dtype=torch.float64, device=u.device)
    tabarr = (C.c_void_p * 6)(*[tabs[k].data_ptr() for k in K.TAB_NAMES])
    stream = torch.cuda.current_stream(u.device).cuda_stream
    with torch.cuda.device(u.device):
        rc = lib.wave3d_cuda_field_error_static(C.byref(pn), C.byref(box), C.byref(lay), tabarr, u.data_ptr(), n,
                                                blocks, out.data_ptr(), stream)
    return rc, out


def _case(N, nranks, rank, seed):
    p = w.Problem(N=N, Lx=1.0, Ly=2.0, Lz=3.0, T=0.1, timesteps=4)
    dec = Decomposition.create(p.N, nranks)
    box, lay = dec.box(rank), dec.layout(rank)
    tabs = K.tables(p, box, DEV)
    g = torch.Generator().manual_seed(seed)
    u = torch.rand(lay.elems, dtype=torch.float64, generator=g).to(DEV)
    return p, box, lay, tabs, u


def test_static_scratch_matches_allocating_entry_point_twice_in_a_row():
    big = _case(96, 1, 0, 11)     # many CTAs: fills a large part of the built-in scratch
    small = _case(40, 8, 5, 12)   # different box, far fewer CTAs: must not see the leftovers
    want_big = K.field_error(*big, 2)       # allocating entry point (scratch from torch)
    want_small = K.field_error(*small, 3)
    rc1, out1 = _static_call(*big, 2)
    rc2, out2 = _static_call(*small, 3)     # second call right behind the first, same stream
    torch.cuda.synchronize()
    assert rc1 == 0 and rc2 == 0, _lib.last_error()
    a1, r1 = out1.tolist()
    a2, r2 = out2.tolist()
    # same kernel, same data, and a maximum does not depend on the order in which the CTAs' partial maxima
    # are combined: both figures must come out bit for bit like the allocating entry point's
    assert (a1, r1) == tuple(want_big)
    assert (a2, r2) == tuple(want_small)


def test_static_scratch_explicit_block_count_and_limit():
    case = _case(64, 1, 0, 13)
    want = K.field_error(*case, 1)
    rc, out = _static_call(*case, 1, blocks=37)   # any CTA count up to the limit gives the same maxima
    torch.cuda.synchronize()
    assert rc == 0, _lib.last_error()
    a, r = out.tolist()
    assert (a, r) == tuple(want)
    rc, _ = _static_call(*case, 1, blocks=1185)   # one more than the built-in scratch can serve
    assert rc != 0
    msg = _lib.last_error()
    assert "1184" in msg and "scratch" in msg
