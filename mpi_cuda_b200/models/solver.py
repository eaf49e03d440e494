"""Public solver API: ``solve(problem, backend=...)`` -> :class:`RunResult`.

One entry point over every implementation the reference shipped as a separate
program (readme.md:5-8, report.pdf §3-§5):

  backend      reference program            implementation here
  ----------   --------------------------   ---------------------------------------
  cpu          wave.cpp (sequential)         native C++ (libwave3d_cpu.so), 1 thread
  cpu-omp      openmpwave.cpp                native C++ + OpenMP, ``threads``
  cpu-ranks    mpi.cpp / mpiomp.cpp          native C++, ranks are threads of one process
  gloo         mpi.cpp / mpiomp.cpp          one process per rank over torch.distributed/gloo
  nccl         mpigpu-1.cu (as written)      pack -> NCCL send/recv -> unpack -> kernel (baseline)
  fused        mpigpu-1.cu (re-designed)     one sm_100a kernel per step, P2P halos in-kernel
  torch        —                             whole-array PyTorch restatement (numerics oracle)
"""
from __future__ import annotations

import time
from typing import Optional, Sequence

import torch
import torch.distributed as dist

from ..utils.report import RunResult
from .problem import WaveProblem

BACKENDS = ("auto", "fused", "nccl", "gloo", "cpu", "cpu-omp", "cpu-ranks", "torch")


def _world() -> int:
    return dist.get_world_size() if dist.is_initialized() else 1


def resolve_dims(p: WaveProblem, decomp: str, world: int):
    from ..parallel.decomp import parse_dims
    return parse_dims(decomp, world, p.unknowns)


def solve(p: WaveProblem, backend: str = "auto", decomp: str = "block", threads: int = 1, every: int = 2,
          ranks: int = 1, verbose: bool = False, check_cfl: bool = True, variant: int = -1,
          zchunk: int = 0, use_graph: bool = True, device: Optional[torch.device] = None,
          pair: bool = True) -> RunResult:
    """Run the whole simulation and return trace + timings.

    ``ranks`` is only used by ``cpu-ranks`` (in-process ranks); for ``gloo`` /
    ``nccl`` / ``fused`` the rank count is the torch.distributed world size.
    ``pair`` (fused backend): temporal blocking, two time steps per pass over memory.
    """
    p.check_cfl(strict=check_cfl)
    if backend == "auto":
        backend = "fused" if torch.cuda.is_available() else ("cpu-omp" if threads > 1 else "cpu")
    if backend not in BACKENDS:
        raise ValueError(f"unknown backend {backend!r}; choose from {BACKENDS}")

    if backend == "torch":
        from . import wave_torch
        t0 = time.perf_counter()
        trace = wave_torch.solve(p, every=every)
        dt = time.perf_counter() - t0
        res = RunResult("torch", p.N, p.L, p.tau, p.K, p.bc, trace=trace, updates=p.updates,
                        phases={"total": dt, "loop": dt})
    elif backend in ("cpu", "cpu-omp", "cpu-ranks"):
        from ..ops import cpu
        nthreads = 1 if backend == "cpu" else max(1, threads)
        dims = resolve_dims(p, decomp, ranks) if backend == "cpu-ranks" else (1, 1, 1)
        trace, phases = cpu.solve(p, threads=nthreads, dims=dims, every=every)
        phases["loop"] = phases["total"] - phases["init"]
        res = RunResult(backend, p.N, p.L, p.tau, p.K, p.bc, world_size=dims[0] * dims[1] * dims[2],
                        decomp=dims, threads=nthreads, trace=trace, phases=phases, updates=p.updates)
    elif backend == "gloo":
        from ..parallel.exchange import CpuBlockOps, ExchangeSolver
        dims = resolve_dims(p, decomp, _world())
        s = ExchangeSolver(p, dims, lambda pp, b: CpuBlockOps(pp, b, threads), every=every, split_shell=True,
                           backend_name="gloo")
        res = s.run()
        res.threads = threads
    elif backend == "nccl":
        from ..parallel.exchange import CudaBlockOps, ExchangeSolver
        dev = device or torch.device("cuda", torch.cuda.current_device())
        dims = resolve_dims(p, decomp, _world())
        s = ExchangeSolver(p, dims, lambda pp, b: CudaBlockOps(pp, b, dev, variant, zchunk), every=every,
                           backend_name="nccl")
        s.run(time_phases=False)          # warm-up (module load, NCCL channels)
        res = s.run(time_phases=True)
    else:  # fused
        from ..parallel.fused import FusedSolver
        dev = device or torch.device("cuda", torch.cuda.current_device())
        dims = resolve_dims(p, decomp, _world())
        s = FusedSolver(p, dims, dev, every=every, variant=variant, zchunk=zchunk, use_graph=use_graph, pair=pair)
        res = s.run()
    rank = dist.get_rank() if dist.is_initialized() else 0
    if verbose and rank == 0:
        for line in res.lines(every):
            print(line)
    return res
