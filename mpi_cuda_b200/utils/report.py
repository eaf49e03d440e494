"""Text and JSON reporting.

Reference parity:
  * per-step line ``Step n, t = ..., Max Error = ..., L2 Error = ...`` printed on
    every second step: report.pdf p.15-16 §4.3 (reproduced byte-for-byte);
  * script banner ``Running test with grid size N^3 [and P processes]``:
    report.pdf p.12;
  * phase timing tables (total / init / compute / boundary / exchange for CPU,
    total / compute / H2D-D2H / MPI for GPU): report.pdf p.16 §4.4;
  * speed-up ``S = T1/T`` and efficiency ``E = S/P``: readme.md:86-100.
"""
from __future__ import annotations

import json
import re
from dataclasses import dataclass, field, asdict
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

STEP_FMT = "Step %d, t = %.6f, Max Error = %.6e, L2 Error = %.6e"
_STEP_RE = re.compile(
    r"Step (\d+), t = ([0-9.]+), Max Error = ([0-9.eE+-]+), L2 Error = ([0-9.eE+-]+)")

# published trace, 512^3, L=1, tau=1e-3 (report.pdf p.15-16 §4.3.1-4.3.2)
GOLDEN_512 = [
    (2, 0.002, 3.967859e-11, 1.406978e-11),
    (4, 0.004, 1.587058e-10, 5.627573e-11),
    (6, 0.006, 3.570526e-10, 1.266079e-10),
    (8, 0.008, 6.346724e-10, 2.250497e-10),
    (10, 0.010, 9.915001e-10, 3.515776e-10),
    (12, 0.012, 1.427449e-09, 5.061618e-10),
    (14, 0.014, 1.942419e-09, 6.887656e-10),
    (16, 0.016, 2.536286e-09, 8.993458e-10),
    (18, 0.018, 3.208909e-09, 1.137852e-09),
    (20, 0.020, 3.960129e-09, 1.404229e-09),
]


def format_step(n: int, t: float, max_err: float, l2_err: float) -> str:
    return STEP_FMT % (n, t, max_err, l2_err)


def parse_step(line: str) -> Optional[Tuple[int, float, float, float]]:
    m = _STEP_RE.search(line)
    if not m:
        return None
    return int(m.group(1)), float(m.group(2)), float(m.group(3)), float(m.group(4))


def banner(N: int, procs: Optional[int] = None) -> str:
    s = f"Running test with grid size {N}^3"
    if procs is not None:
        s += f" and {procs} processes"
    return s


@dataclass
class RunResult:
    """Everything one run produces; JSON-serialisable."""
    backend: str
    N: Tuple[int, int, int]
    L: Tuple[float, float, float]
    tau: float
    K: int
    bc: str
    world_size: int = 1
    decomp: Tuple[int, int, int] = (1, 1, 1)
    threads: int = 1
    trace: List[Tuple[int, float, float, float]] = field(default_factory=list)
    phases: Dict[str, float] = field(default_factory=dict)   # seconds, max over ranks
    updates: int = 0
    extra: Dict[str, object] = field(default_factory=dict)

    @property
    def total_s(self) -> float:
        return float(self.phases.get("total", 0.0))

    @property
    def glups(self) -> float:
        t = self.phases.get("loop", self.total_s)
        return self.updates / t / 1e9 if t > 0 else 0.0

    def lines(self, every: int = 2) -> List[str]:
        return [format_step(*r) for r in self.trace if r[0] % every == 0]

    def to_json(self) -> str:
        d = asdict(self)
        d["glups"] = self.glups
        return json.dumps(d)


CPU_PHASES = ("total", "init", "compute", "boundary", "exchange")
GPU_PHASES = ("total", "init", "compute", "h2d_d2h", "exchange")


def phase_table(res: RunResult) -> str:
    """One-row version of the report's timing tables (report.pdf p.16 §4.4)."""
    names = GPU_PHASES if res.backend in ("fused", "nccl", "gpu") else CPU_PHASES
    note = ""
    if "flag_wait" in res.phases:
        # fused backend: the exchange runs INSIDE the step kernels; its two parts are measured there
        # (%globaltimer, longest face CTA per launch, summed) and are contained in `compute`
        names = names + ("flag_wait", "halo_copy")
        note = "\n(exchange = flag_wait + halo_copy, measured inside the kernels; it overlaps compute and is not added to total)"
    head = " | ".join(f"{n:>10s}" for n in names)
    row = " | ".join(f"{res.phases.get(n, 0.0):10.6f}" if res.backend == "fused" else f"{res.phases.get(n, 0.0):10.4f}"
                     for n in names)
    return f"{head}\n{row}{note}"


def speedup_table(rows: Sequence[Tuple[str, int, float]], t_ref: Optional[float] = None) -> str:
    """rows = (label, P, seconds) -> markdown table with S = T_ref/T and E = S/P.

    Same definitions as the reference's comparison table (readme.md:86-100); the
    first row is the denominator unless ``t_ref`` is given.
    """
    if not rows:
        return ""
    t1 = rows[0][2] if t_ref is None else t_ref
    out = ["| version | P | time s | speed-up | efficiency |", "|---|---|---|---|---|"]
    for label, P, t in rows:
        s = t1 / t if t > 0 else float("nan")
        out.append(f"| {label} | {P} | {t:.4g} | {s:.2f} | {s / P:.2f} |")
    return "\n".join(out)
