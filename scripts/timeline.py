#!/usr/bin/env python
"""Device timeline of one run: per launch start/end (%globaltimer inside the kernels) and the gaps.

usage: python scripts/timeline.py --N 512x512x64 --bc ddp --steps 20 [--pair-variant v] [--no-graph]
(single GPU; under torchrun every rank prints its own table)"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from mpi_cuda_b200 import WaveProblem
from mpi_cuda_b200.parallel.decomp import parse_dims
from mpi_cuda_b200.parallel.fused import FusedSolver
from mpi_cuda_b200.parallel.launch import init_distributed

ap = argparse.ArgumentParser()
ap.add_argument("--N", default="512")
ap.add_argument("--bc", default="ddd")
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--pair-variant", type=int, default=-1)
ap.add_argument("--no-graph", action="store_true")
ap.add_argument("--no-pair", action="store_true")
ap.add_argument("--decomp", default="auto")
a = ap.parse_args()
rank, world, dev = init_distributed("nccl")
Ns = [int(v) for v in a.N.split("x")]
p = WaveProblem(Ns if len(Ns) == 3 else Ns[0], 1.0, 1e-3, a.steps, bc=a.bc)
s = FusedSolver(p, parse_dims(a.decomp, world, p.unknowns), dev, every=2, use_graph=not a.no_graph,
                pair=not a.no_pair, pair_variant=a.pair_variant)
if s.use_graph:
    s.capture()
s.device_timeline()                      # warm
tl = s.device_timeline()
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.init()
ev0.record(); s.run_steps(); ev1.record(); s.barrier()
span = tl[-1][2] - tl[0][1]
busy = sum(e - b for _, b, e in tl)
gaps = [tl[i + 1][1] - tl[i][2] for i in range(len(tl) - 1)]
print(f"rank {rank}: {s.describe().split('||')[-1].strip()}")
for (lv, b, e), g in zip(tl, gaps + [float('nan')]):
    print(f"rank {rank}  level {lv:3d}: start {b:9.2f} us  end {e:9.2f} us  duration {e - b:8.2f}  gap to next {g:6.2f}")
print(f"rank {rank}: launches {len(tl)}, span {span:.1f} us, in-kernel {busy:.1f} us, gaps {sum(gaps):.1f} us "
      f"({100 * sum(gaps) / span:.1f} %), event-timed loop {ev0.elapsed_time(ev1) * 1e3:.1f} us; "
      f"exchange {s.exchange_stats()}")
