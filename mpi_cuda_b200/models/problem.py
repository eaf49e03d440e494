"""Problem definition: grid, boundary conditions, analytic solution, CFL guard.

Reference parity (the reference ships no source; citations are to its documents):
  * PDE, analytic solution and first-kind boundary conditions: report.pdf p.4 §1.
  * Grid ``x_i = i*h``, ``h = L/N``, ``N+1`` nodes per axis, ``t_n = n*tau``:
    report.pdf p.5 §2 (confirmed numerically in SURVEY.md App. A.2).
  * Periodic axes are a north-star extension (BASELINE.json), not in the reference.

Conventions used everywhere in this package
  * ``N`` is the number of *intervals* per axis; a Dirichlet axis has unknowns
    ``i = 1 .. N-1`` (nodes 0 and N are pinned to 0), a periodic axis has unknowns
    ``i = 0 .. N-1`` (node N is node 0).
  * The analytic mode on axis d is ``sin(m_d * pi * x / L_d)`` with ``m_d = 1``
    (Dirichlet) or ``m_d = 2`` (periodic), ``a_t = pi * sqrt(sum m_d^2 / L_d^2)``.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Sequence, Tuple

AXES = ("x", "y", "z")


def _triple(v, cast=float) -> Tuple:
    if isinstance(v, (tuple, list)):
        if len(v) == 1:
            v = v[0]
        else:
            if len(v) != 3:
                raise ValueError(f"expected 1 or 3 values, got {v!r}")
            return tuple(cast(a) for a in v)
    return (cast(v),) * 3


class CFLError(ValueError):
    """Raised when tau violates the leapfrog stability bound."""


@dataclass(frozen=True)
class WaveProblem:
    """Immutable description of one run: ``u_tt = Laplace(u)`` on a box.

    Parameters mirror the reference CLI ``prog N tau K`` (report.pdf p.20 §5.1.3)
    plus the box size ``L`` (the reference ran L = 1 and L = pi, report.pdf p.7-11)
    and per-axis boundary type.
    """

    N: Tuple[int, int, int]
    L: Tuple[float, float, float]
    tau: float
    K: int
    bc: str = "ddd"  # one letter per axis (x, y, z): 'd' Dirichlet, 'p' periodic

    def __init__(self, N, L=1.0, tau=1e-3, K=20, bc="ddd", T=None):
        N3 = _triple(N, int)
        L3 = _triple(L, float)
        if T is not None:
            tau = float(T) / int(K)
        bc = bc.lower()
        if len(bc) == 1:
            bc = bc * 3
        if len(bc) != 3 or any(c not in "dp" for c in bc):
            raise ValueError(f"bc must be 3 letters of d/p, got {bc!r}")
        for n, c in zip(N3, bc):
            if n < (3 if c == "p" else 2):
                raise ValueError(f"N={n} too small for bc {c!r}")
        if min(L3) <= 0 or tau <= 0 or int(K) < 1:
            raise ValueError("L, tau must be positive and K >= 1")
        object.__setattr__(self, "N", N3)
        object.__setattr__(self, "L", L3)
        object.__setattr__(self, "tau", float(tau))
        object.__setattr__(self, "K", int(K))
        object.__setattr__(self, "bc", bc)

    # ---- grid ---------------------------------------------------------------
    @property
    def h(self) -> Tuple[float, float, float]:
        return tuple(l / n for l, n in zip(self.L, self.N))

    @property
    def T(self) -> float:
        return self.tau * self.K

    @property
    def periodic(self) -> Tuple[bool, bool, bool]:
        return tuple(c == "p" for c in self.bc)

    @property
    def unknowns(self) -> Tuple[int, int, int]:
        """Number of updated nodes per axis: N-1 (Dirichlet) or N (periodic)."""
        return tuple(n if p else n - 1 for n, p in zip(self.N, self.periodic))

    @property
    def first_unknown(self) -> Tuple[int, int, int]:
        """Global node index of the first unknown per axis (1 Dirichlet, 0 periodic)."""
        return tuple(0 if p else 1 for p in self.periodic)

    @property
    def n_unknowns(self) -> int:
        a, b, c = self.unknowns
        return a * b * c

    @property
    def n_nodes(self) -> int:
        return (self.N[0] + 1) * (self.N[1] + 1) * (self.N[2] + 1)

    @property
    def updates(self) -> int:
        """Lattice updates of a full run = K * prod(unknowns)."""
        return self.K * self.n_unknowns

    # ---- analytic solution --------------------------------------------------
    @property
    def modes(self) -> Tuple[int, int, int]:
        return tuple(2 if p else 1 for p in self.periodic)

    @property
    def a_t(self) -> float:
        return math.pi * math.sqrt(sum((m / l) ** 2 for m, l in zip(self.modes, self.L)))

    def coeffs(self) -> Tuple[float, float, float]:
        """tau^2 / h_d^2 per axis — the only numbers the stencil kernels need."""
        return tuple(self.tau * self.tau / (hd * hd) for hd in self.h)

    # ---- stability ----------------------------------------------------------
    @property
    def courant(self) -> float:
        """tau * sqrt(sum 1/h_d^2); the leapfrog 7-point scheme needs <= 1."""
        return self.tau * math.sqrt(sum(1.0 / (hd * hd) for hd in self.h))

    def check_cfl(self, strict: bool = True) -> float:
        c = self.courant
        if c > 1.0 and strict:
            raise CFLError(
                f"Courant number {c:.3f} > 1 for N={self.N}, L={self.L}, tau={self.tau}: "
                "the explicit scheme is unstable (use a larger L, a smaller tau, or --no-cfl-check)"
            )
        return c

    def describe(self) -> str:
        return (
            f"grid {self.N[0]}x{self.N[1]}x{self.N[2]} intervals, L={self.L}, tau={self.tau:g}, "
            f"K={self.K}, bc={self.bc}, Courant={self.courant:.3f}"
        )
