"""Domain decomposition planner: 1-D slabs and P x Q x R blocks.

Reference parity: "block partition of the domain between processes" is a stated
requirement of the reference (report.pdf p.4 §1); its MPI programs ran on 1..32
ranks (report.pdf p.9-10) and its CUDA program on 1-2 ranks (report.pdf p.16).
The partition shape is not documented; this planner offers slabs (contiguous
z-planes, two neighbours) and general blocks (up to six neighbours), with the
unknowns of every axis split as evenly as possible (remainder to the low ranks).

Pure Python/integer arithmetic — used identically by the CPU (gloo) backend, the
NCCL baseline and the fused P2P backend, and mirrored by ``split()`` /
``rank_main()`` in csrc/cpu/wave_cpu.cpp.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from ..models.analytic import axis_table
from ..models.problem import WaveProblem


def prime_factors(p: int) -> List[int]:
    f, q = [], 2
    while p > 1:
        if p % q == 0:
            f.append(q)
            p //= q
        else:
            q += 1
    return f


def default_dims(world: int, kind: str = "block") -> Tuple[int, int, int]:
    """Rank grid (Px, Py, Pz).

    ``slab``: (1, 1, P).  ``block``: prime factors of P, largest first, go to z or y,
    whichever is currently smaller (z on ties).  x — the contiguous axis — is only
    split on explicit request, because an x face is an 8-byte-strided column that
    is expensive both to pack and to push over NVLink.
    """
    if kind == "slab":
        return (1, 1, world)
    dims = [1, 1, 1]
    for f in sorted(prime_factors(world), reverse=True):
        t = 2 if dims[2] <= dims[1] else 1
        dims[t] *= f
    return tuple(dims)


def auto_dims(world: int, unknowns: Sequence[int]) -> Tuple[int, int, int]:
    """Measured policy (profiles/scaling.md): z slabs whenever every rank keeps at least 8 planes.

    Every decomposition runs the wide-halo protocol (one launch and one halo handshake per two time
    steps), so what decides is the cost of the faces: z faces leave from registers inside
    the streaming loop, x / y faces are strided copy-outs after it.  On 8 GPUs slabs win at 512^3
    (0.0552 vs 0.0599 ms per step for 1x2x4) and are within 2 % of 1x2x4 blocks at 1024^3 (0.418 vs
    0.409): NVLink 5 is nowhere near a limit, halo volume does not matter.  3-D blocks remain for very
    thin slabs and on request."""
    if world > 1 and unknowns[2] // world >= 8:
        return default_dims(world, "slab")
    return default_dims(world, "block")


def parse_dims(spec: str, world: int, unknowns: Optional[Sequence[int]] = None) -> Tuple[int, int, int]:
    """'slab' | 'block' | 'auto' | 'PxQxR' (x, y, z order)."""
    s = spec.lower()
    if s == "auto":
        return auto_dims(world, unknowns) if unknowns is not None else default_dims(world, "block")
    if s in ("slab", "block"):
        return default_dims(world, s)
    parts = [int(t) for t in s.replace("*", "x").split("x")]
    if len(parts) != 3 or parts[0] * parts[1] * parts[2] != world:
        raise ValueError(f"decomposition {spec!r} does not match world size {world}")
    return tuple(parts)


def split_axis(n: int, parts: int, r: int) -> Tuple[int, int]:
    """(offset, count) of block r when n items are split over parts blocks."""
    q, rem = divmod(n, parts)
    return r * q + min(r, rem), q + (1 if r < rem else 0)


@dataclass(frozen=True)
class RankBlock:
    """What one rank owns and who its neighbours are."""
    rank: int
    coord: Tuple[int, int, int]
    dims: Tuple[int, int, int]
    lo: Tuple[int, int, int]        # global node index of the first owned unknown per axis
    n: Tuple[int, int, int]         # owned unknowns per axis
    nbr: Tuple[Tuple[int, int], Tuple[int, int], Tuple[int, int]]   # [axis][side] rank or -1

    @property
    def n_local(self) -> int:
        return self.n[0] * self.n[1] * self.n[2]

    def has_nbr(self, axis: int, side: int) -> bool:
        return self.nbr[axis][side] >= 0

    def face_elems(self, axis: int) -> int:
        a, c = (axis + 1) % 3, (axis + 2) % 3
        return self.n[a] * self.n[c]

    def halo_bytes_per_step(self) -> int:
        """Bytes this rank must send per time step (8 B per face node)."""
        return 8 * sum(self.face_elems(d) for d in range(3) for s in range(2) if self.nbr[d][s] >= 0)


def rank_of(coord: Sequence[int], dims: Sequence[int]) -> int:
    return (coord[2] * dims[1] + coord[1]) * dims[0] + coord[0]


def coord_of(rank: int, dims: Sequence[int]) -> Tuple[int, int, int]:
    return (rank % dims[0], (rank // dims[0]) % dims[1], rank // (dims[0] * dims[1]))


def plan_rank(p: WaveProblem, dims: Sequence[int], rank: int) -> RankBlock:
    dims = tuple(int(d) for d in dims)
    coord = coord_of(rank, dims)
    lo, n, nbr = [], [], []
    for d in range(3):
        off, cnt = split_axis(p.unknowns[d], dims[d], coord[d])
        if cnt < 1 or (dims[d] > 1 and p.unknowns[d] // dims[d] < 2):
            raise ValueError(f"axis {d}: {p.unknowns[d]} unknowns cannot be split over {dims[d]} ranks")
        lo.append(p.first_unknown[d] + off)
        n.append(cnt)
        sides = []
        for s in (0, 1):
            c = list(coord)
            c[d] += -1 if s == 0 else 1
            if c[d] < 0 or c[d] >= dims[d]:
                if not p.periodic[d]:
                    sides.append(-1)
                    continue
                c[d] %= dims[d]
            sides.append(rank_of(c, dims))
        nbr.append(tuple(sides))
    return RankBlock(rank, coord, dims, tuple(lo), tuple(n), tuple(nbr))


def plan_all(p: WaveProblem, dims: Sequence[int]) -> List[RankBlock]:
    world = dims[0] * dims[1] * dims[2]
    return [plan_rank(p, dims, r) for r in range(world)]


def local_tables(p: WaveProblem, blk: RankBlock, ghosts: int = 1) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Per-axis analytic tables for the local block *including ghosts* (n + 2*ghosts entries).

    Entry l corresponds to global node ``lo - ghosts + l`` (wrapped on periodic axes; 0 for
    nodes beyond a Dirichlet boundary, which only a second ghost layer can reach).
    Values are taken from the global table so that every decomposition sees
    bit-identical numbers.
    """
    out = []
    for d in range(3):
        g = np.arange(blk.lo[d] - ghosts, blk.lo[d] + blk.n[d] + ghosts)
        full = axis_table(p, d)
        if p.periodic[d]:
            t = full[np.mod(g, p.N[d])]
        else:
            inside = (g >= 0) & (g <= p.N[d])
            t = np.where(inside, full[np.clip(g, 0, p.N[d])], 0.0)
        out.append(np.ascontiguousarray(t))
    return tuple(out)


def interior_box(blk: RankBlock) -> Tuple[Tuple[int, int, int], Tuple[int, int, int]]:
    """[lo, hi) in local unknown indices (1-based) of the nodes not adjacent to a
    face that has a neighbour.  Everything else is the shell that must be computed
    and shipped before the neighbours can proceed."""
    lo = tuple(1 + (1 if blk.nbr[d][0] >= 0 else 0) for d in range(3))
    hi = tuple(max(lo[d], blk.n[d] + 1 - (1 if blk.nbr[d][1] >= 0 else 0)) for d in range(3))
    return lo, hi


def shell_boxes(blk: RankBlock) -> List[Tuple[Tuple[int, int, int], Tuple[int, int, int]]]:
    """Disjoint boxes covering the shell: z slabs, then y slabs, then x slabs."""
    ilo, ihi = interior_box(blk)
    lo = [1, 1, 1]
    hi = [blk.n[0] + 1, blk.n[1] + 1, blk.n[2] + 1]
    boxes = []
    for d in (2, 1, 0):
        if ilo[d] > lo[d]:
            h2 = list(hi)
            h2[d] = ilo[d]
            boxes.append((tuple(lo), tuple(h2)))
        if ihi[d] < hi[d]:
            l2 = list(lo)
            l2[d] = ihi[d]
            boxes.append((tuple(l2), tuple(hi)))
        lo[d], hi[d] = ilo[d], ihi[d]
    return [b for b in boxes if all(b[1][k] > b[0][k] for k in range(3))]


def summary(p: WaveProblem, dims: Sequence[int]) -> Dict[str, object]:
    blocks = plan_all(p, dims)
    return {
        "dims": tuple(dims),
        "max_local_unknowns": max(b.n_local for b in blocks),
        "min_local_unknowns": min(b.n_local for b in blocks),
        "max_halo_bytes_per_step": max(b.halo_bytes_per_step() for b in blocks),
        "imbalance": max(b.n_local for b in blocks) * len(blocks) / p.n_unknowns,
    }
