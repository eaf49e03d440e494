#!/usr/bin/env python
"""Collect gpurun_out/suite_*.json into a markdown table (for profiles/)."""
import glob
import json
import os
import re
import sys

rows = []
for f in sorted(glob.glob(os.path.join(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out", "suite_*.json"))):
    m = re.match(r"suite_g(\d+)_(.*)\.json", os.path.basename(f))
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception:
        rows.append((int(m.group(1)), m.group(2), None))
        continue
    rows.append((int(m.group(1)), m.group(2), d))
print("| GPUs | config | decomp | ms/step | wall s | GLUP/s | fraction of roofline | valid | launches/step | clocks MHz (reasons) |")
print("|---|---|---|---|---|---|---|---|---|---|")
for n, name, d in sorted(rows, key=lambda r: (r[1], r[0])):
    if d is None:
        print(f"| {n} | {name} | FAILED | | | | | | | |")
        continue
    c = d["config"]
    print(f"| {n} | {name} ({c['grid']}, K={d['steps']}, {c['backend']}) | {c['parallelism'].replace('spatial ', '')} | "
          f"{d['ms_per_step']:.4f} | {d['wall_s']:.5f} | {d['value']:.1f} | {d['roofline']['fraction_of_roofline']:.3f} | "
          f"{d['validated_vs_oracle']} | {d['gpu_launches'] / d['steps']:.2f} | {d['clocks']['sm_mhz']} ({','.join(d['clocks']['reasons']) or '-'}) |")
