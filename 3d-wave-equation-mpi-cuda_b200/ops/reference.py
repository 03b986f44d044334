"""Plain PyTorch fp64 restatements of every op — the numerics oracle of the tests.

They follow the specification, not the kernels' implementation: whole
arrays, true divisions, per-cell ``abs/|f|``. Results agree with the kernels to the
last bit except where torch has no fused multiply-add (the final ``coef*L + s``),
so the tests compare with a few-ulp tolerance and bit-exactly against the C++
CPU backend, which uses explicit fma()."""
from __future__ import annotations

import math
from typing import Dict, Tuple

import torch

from .. import _lib
from ..utils.spec import Problem


def tables(problem: Problem, box: _lib.Box, device="cpu") -> Dict[str, torch.Tensor]:
    n = problem.native()
    PI = problem.pi
    g = [torch.arange(box.n[a], dtype=torch.float64, device=device) + box.o[a] for a in range(3)]
    sx = torch.sin((((PI + PI) * n.hx) * g[0]) / n.Lx)
    sy = torch.sin(((PI * n.hy) * g[1]) / n.Ly)
    sz = torch.sin(((PI * n.hz) * g[2]) / n.Lz)
    return {"sx": sx, "sy": sy, "sz": sz, "rsx": 1 / sx.abs(), "rsy": 1 / sy.abs(), "rsz": 1 / sz.abs()}


def cos_layer(problem: Problem, n: int) -> float:
    p = problem.native()
    # fma(a_t*n, tau, 2pi): the product a_t*n*tau is far from the rounding boundary of the sum for all
    # practical inputs; math.fma exists from Python 3.13 on
    x = p.a_t * float(n)
    arg = math.fma(x, p.tau, problem.pi + problem.pi) if hasattr(math, "fma") else x * p.tau + (problem.pi + problem.pi)
    return math.cos(arg)


def analytic(problem: Problem, tabs: Dict[str, torch.Tensor], n: int) -> torch.Tensor:
    """[X, Y, Z] analytic solution on the owned points: ((cos*sx)*sy)*sz."""
    c = cos_layer(problem, n)
    return ((c * tabs["sx"])[:, None, None] * tabs["sy"][None, :, None]) * tabs["sz"][None, None, :]


def step(problem: Problem, box: _lib.Box, U1: torch.Tensor, U0: torch.Tensor, tabs: Dict[str, torch.Tensor], n: int,
         first: bool) -> Tuple[torch.Tensor, float, float]:
    """U1: [X+2,Y+2,Z+2] layer n-1 with ghosts; U0: [X,Y,Z] layer n-2 (ignored when first).
    Returns (layer n on the owned points — non-updated cells keep U0's values —, max abs, max rel)."""
    p = problem.native()
    coef = (p.a2 * p.tau) * p.tau
    if first:
        coef = coef * 0.5
    c = U1[1:-1, 1:-1, 1:-1]
    c2 = c + c
    lx = ((U1[2:, 1:-1, 1:-1] + U1[:-2, 1:-1, 1:-1]) - c2) / (p.hx * p.hx)
    ly = ((U1[1:-1, 2:, 1:-1] + U1[1:-1, :-2, 1:-1]) - c2) / (p.hy * p.hy)
    lz = ((U1[1:-1, 1:-1, 2:] + U1[1:-1, 1:-1, :-2]) - c2) / (p.hz * p.hz)
    s = c if first else c2 - U0
    u = coef * ((lx + ly) + lz) + s
    out = U0.clone()
    sl = tuple(slice(box.lo[a], box.hi[a] + 1) for a in range(3))
    out[sl] = u[sl]
    es = tuple(slice(box.elo[a], box.ehi[a] + 1) for a in range(3))
    f = analytic(problem, tabs, n)[es]
    a = (out[es] - f).abs()
    r = a / f.abs()
    mabs = float(torch.nan_to_num(a, nan=0.0).max()) if a.numel() else 0.0
    mrel = float(torch.nan_to_num(r, nan=0.0).max()) if r.numel() else 0.0
    return out, mabs, mrel


def field_error(problem: Problem, box: _lib.Box, U: torch.Tensor, tabs: Dict[str, torch.Tensor], n: int):
    es = tuple(slice(box.elo[a], box.ehi[a] + 1) for a in range(3))
    f = analytic(problem, tabs, n)[es]
    a = (U[es] - f).abs()
    r = a / f.abs()
    return float(torch.nan_to_num(a, nan=0.0).max()), float(torch.nan_to_num(r, nan=0.0).max())


def solve(problem: Problem, device="cpu"):
    """Whole single-box solve in plain PyTorch (small N): returns (max_abs[], max_rel[])."""
    from ..parallel.decomp import Decomposition
    dec = Decomposition.create(problem.N, 1)
    box = dec.box(0)
    tabs = tables(problem, box, device)
    X, Y, Z = box.n
    def pad(U):  # periodic ghost planes with the seam rule, zero ghosts in y/z
        P = torch.zeros(X + 2, Y + 2, Z + 2, dtype=torch.float64, device=device)
        P[1:-1, 1:-1, 1:-1] = U
        P[0, 1:-1, 1:-1] = U[box.send_x[1]]
        P[-1, 1:-1, 1:-1] = U[box.send_x[0]]
        return P
    layers = [analytic(problem, tabs, 0)]
    ea, er = [0.0], [0.0]
    for n in range(1, problem.timesteps + 1):
        first = n == 1
        U0 = torch.zeros_like(layers[0]) if first else layers[-2]
        if n == 2:  # layer 0 carried analytic wall values; later layers have zero walls
            U0 = U0.clone()
            U0[:, 0, :] = 0; U0[:, -1, :] = 0; U0[:, :, 0] = 0; U0[:, :, -1] = 0
        out, a, r = step(problem, box, pad(layers[-1]), U0, tabs, n, first)
        layers.append(out)
        layers = layers[-2:]
        ea.append(a); er.append(r)
    return ea, er
