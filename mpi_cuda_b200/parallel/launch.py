"""Process launch and rendezvous: one process per GPU (or per CPU rank).

Counterpart of the reference's LSF scripts + ``mpirun -np P`` (report.pdf p.12-15,
p.20-26): ``torchrun`` or :func:`spawn` starts the ranks,
``torch.distributed`` (NCCL on GPUs, gloo on CPUs) replaces ``MPI_Init``.
Rendezvous always uses 127.0.0.1 — single node, and container hostnames may not
resolve.
"""
from __future__ import annotations

import os
import socket
import traceback
from typing import Any, Callable, List, Optional, Sequence

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def env_world() -> tuple:
    """(rank, local_rank, world) from the torchrun environment, (0, 0, 1) otherwise."""
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)),
            int(os.environ.get("WORLD_SIZE", 1)))


def init_distributed(backend: Optional[str] = None, device: Optional[torch.device] = None) -> tuple:
    """Initialise the default process group from the torchrun environment.

    Returns (rank, world, device).  With WORLD_SIZE unset/1 nothing is initialised.
    """
    rank, local_rank, world = env_world()
    use_cuda = torch.cuda.is_available() and (backend in (None, "nccl"))
    if device is None:
        device = torch.device("cuda", local_rank) if use_cuda else torch.device("cpu")
    if device.type == "cuda":
        torch.cuda.set_device(device)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        backend = backend or ("nccl" if device.type == "cuda" else "gloo")
        kw = {}
        if backend == "nccl":
            kw["device_id"] = device
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    return rank, world, device


def _worker(rank: int, world: int, port: int, backend: str, fn: Callable, args: tuple, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      LOCAL_RANK=str(rank), WORLD_SIZE=str(world))
    try:
        if backend == "nccl":
            torch.cuda.set_device(rank)
            dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
        else:
            dist.init_process_group("gloo", rank=rank, world_size=world)
        out = fn(rank, world, *args)
        if ret is not None:
            ret[rank] = out
        dist.barrier()
    except Exception:
        if ret is not None:
            ret[rank] = ("__error__", traceback.format_exc())
        raise
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def spawn(fn: Callable, world: int, args: tuple = (), backend: str = "gloo") -> List[Any]:
    """Run ``fn(rank, world, *args)`` on ``world`` fresh processes; returns per-rank results.

    ``fn`` must be importable (module-level).  Used by the tests and by the CLI when
    it is not already running under torchrun.
    """
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = free_port()
        procs = [ctx.Process(target=_worker, args=(r, world, port, backend, fn, args, ret)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join()
        out = [ret.get(r) for r in range(world)]
    bad = [o for o in out if isinstance(o, tuple) and len(o) == 2 and o[0] == "__error__"]
    if bad or any(p.exitcode != 0 for p in procs):
        raise RuntimeError("rank failed:\n" + ("\n".join(b[1] for b in bad) or "non-zero exit"))
    return out
