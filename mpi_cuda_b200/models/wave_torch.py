"""Plain-PyTorch restatement of the difference scheme (CPU or any device).

This is the numerics oracle for every native/CUDA kernel test: same scheme as
report.pdf p.5 §2.1-2.2 (7-point Laplacian, Taylor first step, leapfrog), written
with whole-array torch ops and no custom code.  It is deliberately simple and
slow; it is never on a hot path.
"""
from __future__ import annotations

import math
from typing import List, Optional, Tuple

import torch

from .analytic import axis_tables, time_table
from .problem import WaveProblem


def _lap_term(u: torch.Tensor, p: WaveProblem) -> torch.Tensor:
    """sum_d (tau^2/h_d^2) * (u_{+1} - 2u + u_{-1}) on the full node array.

    ``u`` has shape (Nz+1, Ny+1, Nx+1).  Dirichlet axes: boundary nodes get a
    meaningless value that the caller masks; periodic axes wrap with period N
    (node N duplicates node 0).
    """
    cx, cy, cz = p.coeffs()
    out = torch.zeros_like(u)
    for dim, c, per, n in ((2, cx, p.periodic[0], p.N[0]),
                           (1, cy, p.periodic[1], p.N[1]),
                           (0, cz, p.periodic[2], p.N[2])):
        if per:
            core = u.narrow(dim, 0, n)
            lap = torch.roll(core, 1, dim) - 2.0 * core + torch.roll(core, -1, dim)
            lap = torch.cat([lap, lap.narrow(dim, 0, 1)], dim)
        else:
            lap = torch.zeros_like(u)
            mid = u.narrow(dim, 1, n - 1)
            lap.narrow(dim, 1, n - 1).copy_(u.narrow(dim, 2, n - 1) - 2.0 * mid + u.narrow(dim, 0, n - 1))
        out += c * lap
    return out


def unknown_mask(p: WaveProblem, device="cpu") -> torch.Tensor:
    m = torch.ones((p.N[2] + 1, p.N[1] + 1, p.N[0] + 1), dtype=torch.bool, device=device)
    for dim, per, n in ((2, p.periodic[0], p.N[0]), (1, p.periodic[1], p.N[1]), (0, p.periodic[2], p.N[2])):
        if not per:
            m.narrow(dim, 0, 1).fill_(False)
            m.narrow(dim, n, 1).fill_(False)
    return m


def count_mask(p: WaveProblem, device="cpu") -> torch.Tensor:
    """Nodes that enter the error norms: each unknown exactly once."""
    m = unknown_mask(p, device).clone()
    for dim, per, n in ((2, p.periodic[0], p.N[0]), (1, p.periodic[1], p.N[1]), (0, p.periodic[2], p.N[2])):
        if per:
            m.narrow(dim, n, 1).fill_(False)
    return m


def initial_layers(p: WaveProblem, device="cpu") -> Tuple[torch.Tensor, torch.Tensor]:
    sx, sy, sz = (torch.from_numpy(t).to(device) for t in axis_tables(p))
    u0 = sz[:, None, None] * sy[None, :, None] * sx[None, None, :]
    mask = unknown_mask(p, device)
    u1 = torch.where(mask, u0 + 0.5 * _lap_term(u0, p), torch.zeros_like(u0))
    return u0, u1


def leapfrog(p: WaveProblem, cur: torch.Tensor, prev: torch.Tensor, mask: torch.Tensor) -> torch.Tensor:
    nxt = 2.0 * cur - prev + _lap_term(cur, p)
    return torch.where(mask, nxt, torch.zeros_like(nxt))


def errors(p: WaveProblem, u: torch.Tensor, n: int, cmask: Optional[torch.Tensor] = None,
           phi: Optional[torch.Tensor] = None) -> Tuple[float, float]:
    """(Max Error, L2 Error) of field ``u`` at time level ``n`` (report.pdf p.6 §3.1.2)."""
    if phi is None:
        sx, sy, sz = (torch.from_numpy(t).to(u.device) for t in axis_tables(p))
        phi = sz[:, None, None] * sy[None, :, None] * sx[None, None, :]
    if cmask is None:
        cmask = count_mask(p, u.device)
    e = torch.where(cmask, (u - phi * math.cos(p.a_t * p.tau * n)).abs(), torch.zeros_like(u))
    return float(e.max()), math.sqrt(float((e * e).sum()) / p.n_unknowns)


def solve(p: WaveProblem, device="cpu", every: int = 1, return_field: bool = False):
    """Run K steps; returns [(n, t, max, l2)] at steps divisible by ``every``."""
    u0, u1 = initial_layers(p, device)
    mask = unknown_mask(p, device)
    cmask = count_mask(p, device)
    phi = u0
    prev, cur = u0, u1
    trace: List[Tuple[int, float, float, float]] = []
    for n in range(1, p.K + 1):
        if n > 1:
            prev, cur = cur, leapfrog(p, cur, prev, mask)
        if n % every == 0:
            mx, l2 = errors(p, cur, n, cmask, phi)
            trace.append((n, n * p.tau, mx, l2))
    return (trace, cur) if return_field else trace
