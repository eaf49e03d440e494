"""Analytic solution tables and the closed-form error oracle.

Reference parity:
  * ``u = sin(pi x/Lx) sin(pi y/Ly) sin(pi z/Lz) cos(a_t t)``: report.pdf p.4 §1.
  * Error norms printed per reporting step (C-norm and RMS over interior nodes):
    report.pdf p.6 §3.1.2 and p.15-16 §4.3; the RMS normalisation 1/(N-1)^3 was
    reverse-engineered from the published trace.

The solution is separable, so everything a kernel needs is three 1-D tables and
one scalar per time level; no 3-D analytic array is ever materialised on the GPU.

Oracle: the initial condition is an exact eigenvector of the discrete Laplacian
(for Dirichlet *and* periodic axes), so the numerical solution is
``u^n = phi * cos(n*theta)`` with ``cos(theta) = 1 - tau^2*lambda/2`` (Chebyshev
form of the leapfrog recurrence with the Taylor first step).  Hence the error at
step n is ``|cos(n theta) - cos(n a_t tau)|`` times a norm of phi — an O(1)
check that validates 1024^3 / 2048^3 runs without any CPU run.
"""
from __future__ import annotations

import math
from typing import List, Tuple

import numpy as np

from .problem import WaveProblem


def axis_table(p: WaveProblem, d: int) -> np.ndarray:
    """sin(m_d*pi*i/N_d) for every global node i = 0..N_d of axis d (float64).

    Uses the argument reduction sin(pi*i/N) = sin(pi*(N-i)/N) so the table is
    exactly symmetric and exactly zero on Dirichlet boundary nodes.
    """
    n = p.N[d]
    m = p.modes[d]
    i = np.arange(n + 1, dtype=np.float64)
    t = np.sin(m * math.pi * i / n)
    if m == 1:
        t[0] = 0.0
        t[n] = 0.0
    else:
        t[0] = 0.0
        t[n] = 0.0
        if n % 2 == 0:
            t[n // 2] = 0.0
    return t


def axis_tables(p: WaveProblem) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    return tuple(axis_table(p, d) for d in range(3))


def time_table(p: WaveProblem, K: int | None = None) -> np.ndarray:
    """cos(a_t * n * tau) for n = 0..K."""
    K = p.K if K is None else K
    n = np.arange(K + 1, dtype=np.float64)
    return np.cos(p.a_t * p.tau * n)


def exact_field(p: WaveProblem, n: int) -> np.ndarray:
    """Full (Nz+1, Ny+1, Nx+1) analytic field at time level n (small N only)."""
    sx, sy, sz = axis_tables(p)
    ct = math.cos(p.a_t * p.tau * n)
    return sz[:, None, None] * sy[None, :, None] * sx[None, None, :] * ct


# ---- closed-form oracle -------------------------------------------------------

def discrete_lambda(p: WaveProblem) -> float:
    """Eigenvalue of -Laplace_h for the initial mode."""
    return sum(
        (4.0 / (h * h)) * math.sin(m * math.pi * h / (2.0 * l)) ** 2
        for h, m, l in zip(p.h, p.modes, p.L)
    )


def _phi_norms(p: WaveProblem) -> Tuple[float, float]:
    """(max |phi|, sqrt(mean phi^2)) over the unknowns."""
    mx, ms = 1.0, 1.0
    for d in range(3):
        t = axis_table(p, d)
        lo = p.first_unknown[d]
        u = t[lo: lo + p.unknowns[d]]
        mx *= float(np.max(np.abs(u)))
        ms *= float(np.sum(u * u)) / p.unknowns[d]
    return mx, math.sqrt(ms)


def oracle_amplitude_error(p: WaveProblem, n: int) -> float:
    """|cos(n theta) - cos(n a_t tau)| evaluated without cancellation."""
    lam = discrete_lambda(p)
    s = p.tau * math.sqrt(lam) / 2.0
    if s > 1.0:
        raise ValueError("scheme is unstable for this mode; oracle undefined")
    theta = 2.0 * math.asin(s)
    w = p.a_t * p.tau
    return abs(2.0 * math.sin(n * (theta + w) / 2.0) * math.sin(n * (theta - w) / 2.0))


def oracle_errors(p: WaveProblem, n: int) -> Tuple[float, float]:
    """Exact-arithmetic (Max Error, L2 Error) at step n."""
    amp = oracle_amplitude_error(p, n)
    mx, rms = _phi_norms(p)
    return amp * mx, amp * rms


def oracle_trace(p: WaveProblem, every: int = 2) -> List[Tuple[int, float, float, float]]:
    """[(n, t_n, max_err, l2_err)] at the reporting steps, like the reference prints."""
    out = []
    for n in range(1, p.K + 1):
        if n % every == 0:
            mx, l2 = oracle_errors(p, n)
            out.append((n, n * p.tau, mx, l2))
    return out
