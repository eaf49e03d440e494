"""Roofline arithmetic for the stencil (BASELINE.json north star: "achieved fraction of
roofline, meaning the slower of HBM bytes at peak bandwidth and halo bytes over NVLink").

The reference quotes no roofline; SURVEY.md §6 derives its P100 run at ~13.5 % of that
part's HBM roofline.  Denominators here are the MEASURED ones of this pool
(`MEASURED_PEAKS.json`: STREAM-style copy; peer copy 770 GB/s per direction), with the
recipe's fallback when the file is absent.
"""
from __future__ import annotations

import json
import os
from typing import Dict, Sequence, Tuple

FALLBACK_HBM_GBS = 6650.0      # B200_PROFILING.md fallback
PEER_GBS = 770.0               # measured peer copy per direction per GPU
BYTES_SINGLE_PASS = 24         # read u^n, read u^{n-1}, write u^{n+1}
BYTES_TWO_STEPS_PER_PASS = 16  # read 2 levels + write 2 levels per 2 steps


def measured_peaks(root: str) -> Tuple[float, bool]:
    """(HBM GB/s, True if measured) from MEASURED_PEAKS.json at the repository root."""
    try:
        with open(os.path.join(root, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), True
    except Exception:
        return FALLBACK_HBM_GBS, False


def nodes_with_ghosts(n: Sequence[int]) -> int:
    return (n[0] + 2) * (n[1] + 2) * (n[2] + 2)


def step_roofline(blocks, pair: bool, hbm_gbs: float, peer_gbs: float = PEER_GBS) -> Dict[str, float]:
    """Per-step roofline of the slowest rank.

    blocks: RankBlock list (parallel.decomp.plan_all).  Returns bytes, halo bytes and the
    roofline time in ms = max(HBM time, NVLink time)."""
    bpn = BYTES_TWO_STEPS_PER_PASS if pair else BYTES_SINGLE_PASS
    nodes = max(nodes_with_ghosts(b.n) for b in blocks)
    halo = max(b.halo_bytes_per_step() for b in blocks)
    t_hbm = nodes * bpn / (hbm_gbs * 1e9)
    t_link = halo / (peer_gbs * 1e9)
    return {"nodes_per_gpu": nodes, "bytes_per_node_per_step": bpn, "bytes_per_gpu_per_step": nodes * bpn,
            "halo_bytes_per_gpu_per_step": halo, "roofline_ms_per_step": max(t_hbm, t_link) * 1e3,
            "single_pass_roofline_ms_per_step": max(nodes * BYTES_SINGLE_PASS / (hbm_gbs * 1e9), t_link) * 1e3}
