"""Explicit halo-exchange solver: pack -> send/recv -> unpack -> step.

This is the classical schedule of the reference's distributed programs
(mpi.cpp / mpiomp.cpp / mpigpu-1.cu: "exchange" is a separately timed, additive
phase in report.pdf p.16 §4.4), expressed over ``torch.distributed`` so that one
implementation serves two purposes:

  * backend ``gloo`` + :class:`CpuBlockOps`  — the multi-rank CPU version (MPI
    equivalent) and the "multi-GPU without GPUs" test fixture;
  * backend ``nccl`` + :class:`CudaBlockOps` — the same-box NCCL baseline that the
    fused P2P kernel has to beat (BASELINE.json north star).  It only *calls* NCCL
    for the halo; it is the baseline, not the product.
"""
from __future__ import annotations

import math
import time
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.distributed as dist

from ..models.analytic import time_table
from ..models.problem import WaveProblem
from ..utils.report import RunResult
from .decomp import RankBlock, interior_box, local_tables, plan_rank, shell_boxes


def step_coeffs(n: int) -> Tuple[float, float, float]:
    """(a, b, cscale) for producing level n: Taylor for n = 1, leapfrog after."""
    return (1.0, 0.0, 0.5) if n == 1 else (2.0, -1.0, 1.0)


class CpuBlockOps:
    """Local block on the CPU, native kernels from libwave3d_cpu.so."""

    device = torch.device("cpu")

    def __init__(self, p: WaveProblem, blk: RankBlock, threads: int = 1):
        from ..ops import cpu
        self.cpu, self.p, self.blk, self.threads = cpu, p, blk, threads
        self.b = cpu.make_block(blk.n)
        shape = cpu.block_shape(self.b)
        self.f = [np.zeros(shape), np.zeros(shape)]
        self.tab = local_tables(p, blk)
        self.coef = p.coeffs()
        self.err: Dict[int, Tuple[float, float]] = {}

    def init(self) -> None:
        self.cpu.init(self.f[0], self.b, *self.tab, threads=self.threads)
        self.f[1].fill(0.0)

    def step(self, s: int, a: float, b: float, cscale: float, ct: float, want_err: bool, n: int,
             boxes=None) -> None:
        cur, nxt = self.f[s & 1], self.f[(s & 1) ^ 1]
        c = [cscale * v for v in self.coef]
        if boxes is None:
            self.cpu.step(nxt, cur, self.b, a, b, c, threads=self.threads)
        else:
            for lo, hi in boxes:
                self.cpu.step(nxt, cur, self.b, a, b, c, lo, hi, threads=self.threads)
        if want_err:
            self.err[n] = self.cpu.error(nxt, self.b, *self.tab, ct, threads=self.threads)

    def local_error(self, n: int) -> torch.Tensor:
        return torch.tensor(self.err[n], dtype=torch.float64)

    def pack(self, parity: int, axis: int, side: int) -> torch.Tensor:
        return torch.from_numpy(self.cpu.pack(self.f[parity], self.b, axis, side, ghost=False))

    def recv_buffer(self, axis: int) -> torch.Tensor:
        return torch.empty(self.blk.face_elems(axis), dtype=torch.float64)

    def unpack(self, parity: int, axis: int, side: int, buf: torch.Tensor) -> None:
        self.cpu.unpack(self.f[parity], self.b, axis, side, buf.numpy(), ghost=True)

    def wrap(self, parity: int, axis: int) -> None:
        self.cpu.wrap(self.f[parity], self.b, axis)

    def sync(self) -> None:
        pass

    def field(self, parity: int) -> torch.Tensor:
        """Owned unknowns (no ghosts) as a (nz, ny, nx) tensor."""
        f = self.f[parity]
        return torch.from_numpy(f[1:-1, 1:-1, self.b.xoff + 1: self.b.xoff + 1 + self.blk.n[0]].copy())


class CudaBlockOps:
    """Local block on one B200, kernels from libwave3d_b200.so; no P2P."""

    def __init__(self, p: WaveProblem, blk: RankBlock, device: torch.device, variant: int = -1,
                 zchunk: int = 0, kernel: str = "tma"):
        from ..ops import cuda
        self.cuda, self.p, self.blk, self.kernel = cuda, p, blk, kernel
        self.device = device
        shape = cuda.field_shape(blk.n)
        self.f = [torch.zeros(shape, dtype=torch.float64, device=device) for _ in range(2)]
        self.tab = [torch.from_numpy(t).to(device) for t in local_tables(p, blk)]
        self.flags = torch.zeros(8, dtype=torch.int64, device=device)
        self.status = torch.zeros(4, dtype=torch.int32, device=device)
        self.ws = torch.zeros(cuda.workspace_bytes(), dtype=torch.uint8, device=device)
        self.trace = torch.zeros((p.K + 1, cuda.ERR_SLOTS, 2), dtype=torch.float64, device=device)
        self.plan = cuda.CudaPlan(blk.n, self.f, self.tab, p.coeffs(), self.flags, self.status, self.ws,
                                  variant=variant, zchunk=zchunk)
        self.sendbuf = {(d, s): torch.empty(blk.face_elems(d), dtype=torch.float64, device=device)
                        for d in range(3) for s in range(2) if blk.nbr[d][s] >= 0}

    def init(self) -> None:
        self.plan.init()
        self.trace.zero_()

    def step(self, s: int, a: float, b: float, cscale: float, ct: float, want_err: bool, n: int,
             boxes=None) -> None:
        fn = self.plan.step_tma if self.kernel == "tma" else self.plan.step_box
        if boxes is None:
            fn(s, a, b, cscale, ct, None, None, self.trace[n, 0] if want_err else None, 0)
        else:
            for q, (lo, hi) in enumerate(boxes):
                # thin shell boxes go through the per-node kernel
                f2 = fn if q == 0 else self.plan.step_box
                f2(s, a, b, cscale, ct, lo, hi, self.trace[n, q] if want_err else None, q)

    def local_error(self, n: int) -> torch.Tensor:
        t = self.trace[n]
        return torch.stack([t[:, 0].max(), t[:, 1].sum()])

    def pack(self, parity: int, axis: int, side: int) -> torch.Tensor:
        buf = self.sendbuf[(axis, side)]
        self.plan.pack_face(parity, axis, side, buf, ghost=False)
        return buf

    def recv_buffer(self, axis: int) -> torch.Tensor:
        return torch.empty(self.blk.face_elems(axis), dtype=torch.float64, device=self.device)

    def unpack(self, parity: int, axis: int, side: int, buf: torch.Tensor) -> None:
        self.plan.unpack_face(parity, axis, side, buf, ghost=True)

    def wrap(self, parity: int, axis: int) -> None:
        self.plan.wrap_axis(parity, axis)

    def sync(self) -> None:
        torch.cuda.synchronize(self.device)

    def field(self, parity: int) -> torch.Tensor:
        xo = self.cuda.XOFF + 1
        return self.f[parity][1:-1, 1:-1, xo: xo + self.blk.n[0]].contiguous()


class ExchangeSolver:
    """pack -> isend/irecv -> unpack -> step, every time step, on every rank."""

    def __init__(self, p: WaveProblem, dims: Sequence[int], ops_factory, group=None, every: int = 2,
                 split_shell: bool = False, backend_name: str = "exchange"):
        self.p, self.dims, self.every = p, tuple(dims), every
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        if self.world != dims[0] * dims[1] * dims[2]:
            raise ValueError(f"dims {dims} need {dims[0] * dims[1] * dims[2]} ranks, world size is {self.world}")
        self.blk = plan_rank(p, dims, self.rank)
        self.ops = ops_factory(p, self.blk)
        self.split_shell = split_shell
        self.backend_name = backend_name
        self.recvbuf = {(d, s): self.ops.recv_buffer(d) for d in range(3) for s in range(2)
                        if self.blk.nbr[d][s] >= 0 and self.blk.nbr[d][s] != self.rank}
        self.ct = time_table(p)
        self.graph = None                 # CUDA graph of the whole K-step loop (GPU ops only, see capture())
        self._err_gather = None

    # -- the baseline's fair shot: the whole loop in one CUDA graph -------------
    def _loop_body(self) -> None:
        """K steps on the current stream, no host synchronisation: exchange -> step -> on reporting
        steps ONE all-gather of this rank's (max, sum) pair (combined on the host after the loop)."""
        p, ops = self.p, self.ops
        for n in range(1, p.K + 1):
            s = n - 1
            a, b, cs = step_coeffs(n)
            want = n % self.every == 0
            self.exchange(s & 1)
            ops.step(s, a, b, cs, float(self.ct[n]), want, n, None)
            if want:
                e = ops.local_error(n)
                if self.world > 1:
                    dist.all_gather_into_tensor(self._err_gather[n], e, group=self.group)
                else:
                    self._err_gather[n].copy_(e)

    def capture(self) -> bool:
        """Capture the K-step loop (pack, NCCL send/recv, unpack, step, error all-gather) in a CUDA graph.

        Returns False (and leaves the solver in eager mode) if NCCL refuses to be captured."""
        if self.graph is not None:
            return True
        dev = self.ops.device
        self._err_gather = torch.zeros((self.p.K + 1, 2 * self.world), dtype=torch.float64, device=dev)
        self.ops.init()
        l0 = self.ops.plan.launches
        self._loop_body()                    # warm-up outside capture (lazy module load, NCCL channels)
        self.launches_per_run = self.ops.plan.launches - l0
        torch.cuda.synchronize(dev)
        if self.world > 1:
            dist.barrier(self.group)
        try:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._loop_body()
            self.graph = g
        except Exception as e:               # pragma: no cover - depends on the NCCL build
            self.graph = None
            self.capture_error = repr(e)
            torch.cuda.synchronize(dev)
            return False
        torch.cuda.synchronize(dev)
        return True

    def run_graph(self) -> Tuple[float, List[Tuple[int, float, float, float]]]:
        """One whole run by graph replay; returns (loop seconds on the device, max over ranks; trace)."""
        p, dev = self.p, self.ops.device
        self.ops.init()
        torch.cuda.synchronize(dev)
        if self.world > 1:
            dist.barrier(self.group)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        self.graph.replay()
        ev1.record()
        torch.cuda.synchronize(dev)
        t = torch.tensor([ev0.elapsed_time(ev1) / 1e3], dtype=torch.float64, device=dev)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        g = self._err_gather.cpu().view(p.K + 1, self.world, 2)
        trace = [(n, n * p.tau, float(g[n, :, 0].max()), math.sqrt(float(g[n, :, 1].sum()) / p.n_unknowns))
                 for n in range(1, p.K + 1) if n % self.every == 0]
        return float(t.item()), trace

    # -- halo exchange of the array with the given parity ---------------------
    def exchange(self, parity: int) -> None:
        blk, ops = self.blk, self.ops
        reqs: List[dist.P2POp] = []
        for d in range(3):
            if blk.nbr[d][0] == self.rank:          # periodic axis owned by one rank
                ops.wrap(parity, d)
                continue
            for s in range(2):
                peer = blk.nbr[d][s]
                if peer < 0:
                    continue
                reqs.append(dist.P2POp(dist.isend, ops.pack(parity, d, s), peer, self.group, tag=d * 2 + (1 - s)))
                reqs.append(dist.P2POp(dist.irecv, self.recvbuf[(d, s)], peer, self.group, tag=d * 2 + s))
        if reqs:
            if ops.device.type == "cpu":
                # gloo: plain isend/irecv with tags (batch_isend_irecv ignores tags on some backends)
                works = [op.op(op.tensor, op.peer, group=op.group, tag=op.tag) for op in reqs]
            else:
                # NCCL matches sends and recvs of a peer pair in issue order; tags are ignored.
                # Send tags carry the *receiver's* face id, so ordering both lists by
                # (peer, tag) makes the two sides agree even when one peer is both my
                # low and my high neighbour (periodic axis split over two ranks).
                sends = sorted((o for o in reqs if o.op is dist.isend), key=lambda o: (o.peer, o.tag))
                recvs = sorted((o for o in reqs if o.op is dist.irecv), key=lambda o: (o.peer, o.tag))
                works = dist.batch_isend_irecv(sends + recvs)
            for w in works:
                w.wait()
            for d in range(3):
                for s in range(2):
                    if (d, s) in self.recvbuf:
                        ops.unpack(parity, d, s, self.recvbuf[(d, s)])

    def run(self, verbose: bool = False, time_phases: bool = True) -> RunResult:
        p, ops = self.p, self.ops
        ph = {"total": 0.0, "init": 0.0, "compute": 0.0, "boundary": 0.0, "exchange": 0.0, "loop": 0.0}
        boxes = None
        if self.split_shell:
            boxes = [interior_box(self.blk)] + shell_boxes(self.blk)
        trace: List[Tuple[int, float, float, float]] = []
        pending: List[Tuple[int, torch.Tensor]] = []

        def tick():
            if time_phases:
                ops.sync()
            return time.perf_counter()

        t_start = tick()
        ops.init()
        if self.world > 1:
            dist.barrier(self.group)
        t0 = tick()
        ph["init"] = t0 - t_start
        on_gpu = ops.device.type == "cuda"
        if on_gpu:                          # the loop time is taken on the device (CUDA events)
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
        for n in range(1, p.K + 1):
            s = n - 1
            a, b, cs = step_coeffs(n)
            want = n % self.every == 0
            t1 = tick()
            self.exchange(s & 1)
            t2 = tick()
            ops.step(s, a, b, cs, float(self.ct[n]), want, n, boxes)
            t3 = tick()
            ph["exchange"] += t2 - t1
            ph["compute"] += t3 - t2
            if want:
                e = ops.local_error(n)
                if self.world > 1:
                    mx, s2 = e[0:1].clone(), e[1:2].clone()
                    dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=self.group)
                    dist.all_reduce(s2, op=dist.ReduceOp.SUM, group=self.group)
                    e = torch.cat([mx, s2])
                pending.append((n, e))
                ph["exchange"] += tick() - t3
        if on_gpu:
            ev1.record()
        ops.sync()
        t_end = time.perf_counter()
        ph["loop"] = ev0.elapsed_time(ev1) / 1e3 if on_gpu else t_end - t0
        ph["total"] = t_end - t_start
        for n, e in pending:
            mx, s2 = (float(v) for v in e.cpu())
            trace.append((n, n * p.tau, mx, math.sqrt(s2 / p.n_unknowns)))
        if self.world > 1:
            t = torch.tensor([ph[k] for k in sorted(ph)], dtype=torch.float64, device=ops.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
            ph = dict(zip(sorted(ph), (float(v) for v in t.cpu())))
        res = RunResult(backend=self.backend_name, N=p.N, L=p.L, tau=p.tau, K=p.K, bc=p.bc,
                        world_size=self.world, decomp=self.dims, trace=trace, phases=ph, updates=p.updates)
        if verbose and self.rank == 0:
            for line in res.lines(self.every):
                print(line)
        return res
