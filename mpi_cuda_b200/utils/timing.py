"""Phase timers and step traces.

Reference parity: the reference reports hand-placed wall-clock phase timers summed
over the run (report.pdf p.16 §4.4: total / init / compute / boundary / exchange on
CPU, total / compute / H2D-D2H / MPI on GPU).  Here device phases are timed with
CUDA events on the launching stream (max over ranks is taken by the caller), and a
per-step trace can be written in Chrome trace-event format for chrome://tracing /
Perfetto.
"""
from __future__ import annotations

import json
import time
from contextlib import contextmanager
from typing import Dict, List, Optional

import torch


class PhaseTimer:
    """Accumulates named phases; CUDA events when a device is given, perf_counter otherwise."""

    def __init__(self, device: Optional[torch.device] = None):
        self.device = device if (device is not None and device.type == "cuda") else None
        self.totals: Dict[str, float] = {}
        self._pending: List = []

    @contextmanager
    def phase(self, name: str):
        if self.device is None:
            t0 = time.perf_counter()
            yield
            self.totals[name] = self.totals.get(name, 0.0) + time.perf_counter() - t0
        else:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            yield
            e1.record()
            self._pending.append((name, e0, e1))

    def result(self) -> Dict[str, float]:
        if self.device is not None:
            torch.cuda.synchronize(self.device)
            for name, e0, e1 in self._pending:
                self.totals[name] = self.totals.get(name, 0.0) + e0.elapsed_time(e1) / 1e3
            self._pending.clear()
        return dict(self.totals)


def write_chrome_trace(path: str, step_ms: List[float], rank: int = 0, name: str = "fused_step",
                       device_spans=None) -> None:
    """step_ms[i] = device time of step i+1 (CUDA events); written as back-to-back complete events on
    track 0.  ``device_spans`` = [(input level, start_us, end_us)] from ``FusedSolver.device_timeline``
    (stamps taken inside the kernels) go on track 1: the distance between consecutive spans is the
    launch gap, which event timing cannot see."""
    t = 0.0
    ev = []
    for i, ms in enumerate(step_ms):
        ev.append({"name": f"{name} {i + 1}", "ph": "X", "ts": t * 1e3, "dur": ms * 1e3, "pid": rank, "tid": 0})
        t += ms
    for lv, b, e in (device_spans or []):
        ev.append({"name": f"launch from level {lv}", "ph": "X", "ts": b, "dur": e - b, "pid": rank, "tid": 1})
    with open(path, "w") as fh:
        json.dump({"traceEvents": ev, "displayTimeUnit": "ms"}, fh)
