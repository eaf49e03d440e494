"""Checkpoint / restart of a running simulation (one file per rank).

The reference has no checkpointing (its runs are K = 20 steps); this
is the optional dump/restart of the two live time levels that the survey plans.
A checkpoint taken after level ``n`` holds both local arrays *including ghost
layers*, so a restart needs no exchange before its first step.
"""
from __future__ import annotations

import os
from typing import Dict

import torch


def rank_path(path: str, rank: int) -> str:
    return f"{path}.rank{rank}.pt"


def save_state(path: str, rank: int, level: int, fields, meta: Dict) -> str:
    out = rank_path(path, rank)
    tmp = out + ".tmp"
    torch.save({"level": int(level), "fields": [f.detach().cpu() for f in fields], "meta": meta}, tmp)
    os.replace(tmp, out)           # atomic: a crash never leaves a half-written checkpoint
    return out


def load_state(path: str, rank: int) -> Dict:
    return torch.load(rank_path(path, rank), map_location="cpu", weights_only=False)
