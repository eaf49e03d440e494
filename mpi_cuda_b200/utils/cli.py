"""Command line: ``python -m mpi_cuda_b200 N tau K [threads]`` or ``... N L T K``.

Reference parity: every invocation shown in the reference is positional
``prog N tau K [threads]`` (report.pdf p.20 §5.1.3, p.21 §5.2.3, p.22 §5.3.3,
p.15 §4.2.4); BASELINE.json additionally asks for ``N L T K``.  Both are accepted:

  * 3 positionals                     -> N tau K
  * 4 positionals, 3rd is an integer
    literal and 2nd <= 0.1            -> N tau K threads   (reference style)
  * 4 positionals otherwise           -> N L T K           (tau = T / K)

``--style ref|nltk`` forces one reading.  Output: the reference's banner and
``Step ...`` lines (report.pdf p.12, p.15-16), then a timing table and, with
``--json``, one machine-readable line.
Multi-GPU: launch under torchrun (one process per GPU); multi-rank CPU:
``--backend gloo`` under torchrun or ``--backend cpu-ranks --ranks P`` in-process.
"""
from __future__ import annotations

import argparse
import math
import sys
from typing import List, Optional, Tuple

from ..models.problem import WaveProblem


def _is_int_literal(s: str) -> bool:
    return s.lstrip("+-").isdigit()


def _num(s: str) -> float:
    s = s.strip().lower()
    if s in ("pi", "π"):
        return math.pi
    if s.endswith("pi"):
        return float(s[:-2] or 1) * math.pi
    return float(s)


def interpret_positionals(pos: List[str], style: str = "auto") -> dict:
    """-> dict(N, tau, K, L or None, threads or None)."""
    if len(pos) == 3:
        return dict(N=int(pos[0]), tau=_num(pos[1]), K=int(pos[2]), L=None, threads=None)
    if len(pos) != 4:
        raise SystemExit("expected 'N tau K [threads]' or 'N L T K'")
    if style == "auto":
        style = "ref" if (_is_int_literal(pos[2]) and _num(pos[1]) <= 0.1) else "nltk"
    if style == "ref":
        return dict(N=int(pos[0]), tau=_num(pos[1]), K=int(pos[2]), L=None, threads=int(pos[3]))
    K = int(pos[3])
    return dict(N=int(pos[0]), tau=_num(pos[2]) / K, K=K, L=_num(pos[1]), threads=None)


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(prog="mpi_cuda_b200", description=__doc__,
                                 formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("positional", nargs="+", help="N tau K [threads]  |  N L T K")
    ap.add_argument("--style", choices=("auto", "ref", "nltk"), default="auto")
    ap.add_argument("--L", nargs="+", default=None, help="box size: one value or Lx Ly Lz ('pi' accepted)")
    ap.add_argument("--bc", default="ddd", help="per-axis boundary: d (Dirichlet) or p (periodic), e.g. ddp")
    ap.add_argument("--backend", default="auto",
                    choices=("auto", "fused", "nccl", "gloo", "cpu", "cpu-omp", "cpu-ranks", "torch"))
    ap.add_argument("--decomp", default="auto", help="auto (measured policy: z slabs) | slab | block | PxQxR")
    ap.add_argument("--ranks", type=int, default=1, help="in-process ranks for --backend cpu-ranks")
    ap.add_argument("--threads", type=int, default=None, help="OpenMP threads per rank")
    ap.add_argument("--report-every", type=int, default=2)
    ap.add_argument("--no-cfl-check", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch step kernels one by one instead of a CUDA graph")
    ap.add_argument("--no-pair", action="store_true", help="fused backend: one pass per time step (no temporal blocking)")
    ap.add_argument("--variant", type=int, default=-1, help="TMA tile variant (see DESIGN.md)")
    ap.add_argument("--zchunk", type=int, default=0)
    ap.add_argument("--json", action="store_true")
    ap.add_argument("--quiet", action="store_true")
    ap.add_argument("--oracle", action="store_true", help="also print the closed-form expected errors")
    ap.add_argument("--checkpoint", metavar="PATH", help="fused backend: dump the state after --checkpoint-at steps")
    ap.add_argument("--checkpoint-at", type=int, default=0)
    ap.add_argument("--restart", metavar="PATH", help="fused backend: continue from a dump written by --checkpoint")
    ap.add_argument("--trace", metavar="FILE", help="fused backend: per-step device times as a Chrome trace")
    return ap


def parse(argv: Optional[List[str]] = None) -> Tuple[WaveProblem, argparse.Namespace]:
    args = build_parser().parse_args(argv)
    f = interpret_positionals(args.positional, args.style)
    L = f["L"] if f["L"] is not None else 1.0
    if args.L is not None:
        L = [_num(v) for v in args.L]
    threads = args.threads if args.threads is not None else (f["threads"] or 1)
    args.threads = threads
    p = WaveProblem(f["N"], L, f["tau"], f["K"], bc=args.bc)
    return p, args


def _fused_session(p: WaveProblem, args, rank: int) -> int:
    """--checkpoint / --restart / --trace: drive the fused solver step ranges directly."""
    import torch
    import torch.distributed as dist
    from ..parallel.decomp import parse_dims
    from ..parallel.fused import FusedSolver
    from ..utils.report import format_step
    from ..utils.timing import write_chrome_trace
    p.check_cfl(strict=not args.no_cfl_check)
    world = dist.get_world_size() if dist.is_initialized() else 1
    s = FusedSolver(p, parse_dims(args.decomp, world, p.unknowns), every=args.report_every, variant=args.variant,
                    zchunk=args.zchunk, use_graph=False, pair=not args.no_pair)
    if args.trace:
        ms = s.profile_steps()
        tl = s.device_timeline()           # %globaltimer stamps from inside the kernels: per-launch spans and gaps
        xs = s.exchange_stats()
        if rank == 0:
            write_chrome_trace(args.trace, ms, rank, device_spans=tl)
            print(f"wrote {args.trace}: {len(ms)} steps, mean {sum(ms) / len(ms):.4f} ms/step; "
                  f"{len(tl)} launches, in-kernel flag wait {xs['flag_wait_s'] * 1e6:.1f} us, "
                  f"halo copy-out {xs['halo_copy_s'] * 1e6:.1f} us")
    else:
        first = 1
        if args.restart:
            first = s.load_checkpoint(args.restart) + 1
        else:
            s.init()
        last = args.checkpoint_at if (args.checkpoint and 0 < args.checkpoint_at < p.K) else p.K
        s.run_range(first, last)
        if args.checkpoint and last < p.K:
            out = s.save_checkpoint(args.checkpoint, last)
            if rank == 0:
                print(f"checkpoint after step {last}: {out}")
        s.barrier()
        s.check_status()
    if rank == 0 and not args.quiet:
        for r in s.gather_trace():
            if r[2] > 0 or r[3] > 0:
                print(format_step(*r))
    return 0


def main(argv: Optional[List[str]] = None) -> int:
    import torch
    import torch.distributed as dist
    from ..models.solver import solve
    from ..parallel.launch import env_world, init_distributed
    from ..utils.report import banner, phase_table

    p, args = parse(argv)
    rank, _, world = env_world()
    if world > 1:
        backend = "gloo" if args.backend in ("gloo", "cpu", "cpu-omp", "cpu-ranks", "torch") else None
        init_distributed(backend)
    elif args.backend in ("auto", "fused", "nccl") and torch.cuda.is_available():
        torch.cuda.set_device(0)
    procs = world if world > 1 else (args.ranks if args.backend == "cpu-ranks" else None)
    if rank == 0 and not args.quiet:
        print(banner(p.N[0], procs))
        if p.courant > 1.0:
            print(f"warning: Courant number {p.courant:.3f} > 1", file=sys.stderr)
    try:
        if args.checkpoint or args.restart or args.trace:
            return _fused_session(p, args, rank)
        res = solve(p, backend=args.backend, decomp=args.decomp, threads=args.threads,
                    every=args.report_every, ranks=args.ranks, verbose=not args.quiet,
                    check_cfl=not args.no_cfl_check, variant=args.variant, zchunk=args.zchunk,
                    use_graph=not args.no_graph, pair=not args.no_pair)
        if rank == 0:
            if not args.quiet:
                print(phase_table(res))
                print(f"backend {res.backend}, ranks {res.world_size} {res.decomp}, threads {res.threads}, "
                      f"{res.glups:.3f} GLUP/s")
            if args.oracle:
                from ..models.analytic import oracle_trace
                from ..utils.report import format_step
                for r in oracle_trace(p, args.report_every):
                    print("oracle " + format_step(*r))
            if args.json:
                print(res.to_json())
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()
    return 0
