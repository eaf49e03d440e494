"""Fused multi-GPU solver: the halo exchange and the error reduction live inside the step kernels.

This is the product path (BASELINE.json north star).  Where the reference's
MPI+CUDA program staged halos D2H -> MPI_Sendrecv -> H2D every step (separate
"H2D/D2H" and "MPI" columns in report.pdf p.16 §4.4) and reduced the error with
MPI, here:

  * every rank's field arrays, its flag words and its error table live in
    symmetric memory (``torch.distributed._symmetric_memory``), so a kernel on
    GPU A can store to GPU B's arrays over NVLink 5 / NVSwitch;
  * the step kernels number the CTAs that touch a face with a neighbour first;
    those acquire the neighbour's flag before loading, and after their tile is
    written they copy its face layers straight into the neighbour's ghost layers
    and release a system-scope flag — no host, no NCCL inside the time loop;
  * the error max / sum-of-squares is an epilogue of the same kernels, and the
    last CTA stores the pair into every rank's table, so after the loop each rank
    already holds all partials (no collective at all);
  * by default two time steps share one pass over memory (``pair=True``:
    ``w3d_step_pair`` + ``w3d_step_fixup``); ``pair=False`` runs
    one ``w3d_step_fused`` launch per step;
  * the K steps of a run are captured in one CUDA graph.

With world size 1 the same code runs on plain device tensors (a periodic axis
wraps onto the rank itself through the same peer mechanism).
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from ..models.analytic import time_table
from ..models.problem import WaveProblem
from ..ops import cuda
from ..utils.report import RunResult
from .decomp import RankBlock, local_tables, plan_all, plan_rank
from .exchange import step_coeffs

_FLAG_BYTES = 256     # 6 flag words + padding, at the head of the symmetric buffer
_SLOTS = 2            # error-table entries per rank and level (step kernel, fix-up pass)


def _field_bytes(n: Sequence[int]) -> int:
    a, b, c = cuda.field_shape(n)
    return a * b * c * 8


class FusedSolver:
    def __init__(self, p: WaveProblem, dims: Sequence[int] = (1, 1, 1), device: Optional[torch.device] = None,
                 group=None, every: int = 2, variant: int = -1, zchunk: int = 0, use_graph: bool = True,
                 spin_timeout_ms: int = 0, pair: bool = False, pair_variant: int = -1):
        self._launches_per_run = p.K
        # pair mode with an even K never reads array 1 before overwriting it
        self._zero_second = not (pair and p.K % 2 == 0)
        self.p, self.dims, self.every = p, tuple(int(d) for d in dims), every
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        if self.world != self.dims[0] * self.dims[1] * self.dims[2]:
            raise ValueError(f"dims {self.dims} do not match world size {self.world}")
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.use_graph = use_graph
        blocks = plan_all(p, self.dims)
        self.blk = blocks[self.rank]
        blk = self.blk
        fbytes = max(_field_bytes(b.n) for b in blocks)
        fbytes = (fbytes + 255) // 256 * 256
        # error table: one row per rank, (K+1) x (max, sum sq); every rank holds a full copy
        # that the other ranks' kernels fill over NVLink
        tbytes = (self.world * (p.K + 1) * _SLOTS * 16 + 255) // 256 * 256
        f0 = _FLAG_BYTES + tbytes                     # offset of the first field array
        self.pair = bool(pair)
        narr = 4 if self.pair else 2                  # temporal blocking needs two more arrays
        total = f0 + narr * fbytes
        # ---- storage: symmetric when there are real peers ----
        self.symm = None
        if self.world > 1:
            import torch.distributed._symmetric_memory as symm_mem
            buf = symm_mem.empty(total, dtype=torch.uint8, device=self.device)
            self.symm = symm_mem.rendezvous(buf, group if group is not None else dist.group.WORLD)
            base_ptrs = [int(x) for x in self.symm.buffer_ptrs]
        else:
            buf = torch.empty(total, dtype=torch.uint8, device=self.device)
            base_ptrs = [buf.data_ptr()]
        buf.zero_()
        self.buf = buf
        shape = cuda.field_shape(blk.n)
        nel = shape[0] * shape[1] * shape[2]
        self.flags = buf[:_FLAG_BYTES].view(torch.int64)
        self.f = [buf[f0 + q * fbytes: f0 + q * fbytes + nel * 8].view(torch.float64).view(shape)
                  for q in range(2)]
        self.aux = [buf[f0 + q * fbytes: f0 + q * fbytes + nel * 8].view(torch.float64).view(shape)
                    for q in range(2, narr)]
        # (world, K+1, slots, 2): slot 0/1 = the step kernels, slot 1 also = the fix-up pass of a pair
        self.table = buf[_FLAG_BYTES: _FLAG_BYTES + self.world * (p.K + 1) * _SLOTS * 16].view(torch.float64).view(
            self.world, p.K + 1, _SLOTS, 2)
        self.tab = [torch.from_numpy(t).to(self.device) for t in local_tables(p, blk)]
        self.status = torch.zeros(4, dtype=torch.int32, device=self.device)
        self.ws = torch.zeros(cuda.workspace_bytes(), dtype=torch.uint8, device=self.device)
        self.trace = self.table[self.rank, :, 0]      # my row: the kernel writes it here and in every peer's copy
        self.trace_fix = self.table[self.rank, :, 1]  # fix-up pass of the two-step scheme
        # ---- peers ----
        peers: Dict[int, cuda.PeerSpec] = {}
        peer_aux: Dict[int, Tuple[int, int]] = {}
        for d in range(3):
            for s in range(2):
                r = blk.nbr[d][s]
                if r < 0:
                    continue
                pb = blocks[r]
                pshape = cuda.field_shape(pb.n)
                # my low edge fills the peer's HIGH ghost (index n+1), my high edge its LOW ghost (0)
                ghost = pb.n[d] + 1 if s == 0 else 0
                base = base_ptrs[r] if self.world > 1 else base_ptrs[0]
                peers[d * 2 + s] = cuda.PeerSpec(
                    field_ptrs=(base + f0, base + f0 + fbytes), ghost_index=ghost,
                    pitch=pshape[2], plane=pshape[1] * pshape[2],
                    flag_ptr=base + 8 * (d * 2 + (1 - s)))
                peer_aux[d * 2 + s] = (base + f0 + 2 * fbytes, base + f0 + 3 * fbytes)
        self.plan = cuda.CudaPlan(blk.n, self.f, self.tab, p.coeffs(), self.flags, self.status, self.ws,
                                  peers=peers, variant=variant, zchunk=zchunk, spin_timeout_ms=spin_timeout_ms,
                                  err_table_ptr=base_ptrs[self.rank if self.world > 1 else 0] + _FLAG_BYTES,
                                  err_table_peer_ptrs=[base_ptrs[r] + _FLAG_BYTES for r in range(self.world)
                                                       if r != self.rank] if self.world > 1 else [])
        self.ct = time_table(p)
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.kernel_launches = 0
        # ---- temporal blocking (two steps per pass) ----
        self.final_index = None          # array index (0..3) holding level K after run_steps in pair mode
        if self.pair:
            self.plan.enable_pairs(self.aux[0], self.aux[1], pair_variant, zchunk, peer_aux)

    # ------------------------------------------------------------------
    def describe(self) -> str:
        return self.plan.describe()

    def barrier(self) -> None:
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            dist.barrier(self.group)

    def init(self) -> None:
        """Layer 0 everywhere (ghosts included), zero second array, flags reset.

        All ranks must have finished this before any rank's first step may push
        into their ghosts, hence the barrier."""
        self.plan.init(zero_second=self._zero_second)
        self.table[self.rank].zero_()                    # only my own row: peers clear theirs
        self.barrier()

    def step(self, n: int, want_err: bool = True) -> None:
        """Produce time level n (n = 1 is the Taylor step)."""
        a, b, cs = step_coeffs(n)
        if n == 1 and not self._zero_second:
            # init() skipped the zero fill of array 1 (even-K pair runs never read it); this
            # single Taylor step does read it (times b = 0), so make sure it holds finite zeros
            self.f[1].zero_()
        self.plan.step_fused(n - 1, a, b, cs, float(self.ct[n]), self.trace[n] if want_err else None)
        self.kernel_launches += 1

    def _enqueue_steps(self, first: int, last: int) -> None:
        if self.pair and first == 1:
            self._enqueue_pairs(last)
            return
        for n in range(first, last + 1):
            self.step(n, n % self.every == 0)

    def _enqueue_pairs(self, last: int) -> None:
        """Levels 1..last with two steps per pass (plus one single step if `last` is odd).

        Array roles rotate over the four arrays: (cur, prev) -> (out1, out2) = (0,1)->(2,3),
        (3,2)->(0,1), (1,0)->(2,3), ...  With neighbours (or periodic axes) every pair is two
        launches: the pair pass (which also ships the level-(n+1) faces) and the fix-up pass
        that finishes level n+2 next to those faces and ships them."""
        n = self.pair_begin(last)
        while n < last:
            n = self.pair_advance(n)

    def pair_begin(self, last: int) -> int:
        """Start a pair-mode run of `last` steps; returns the level reached (0, or 1 if odd)."""
        self._cur, self._prev = 0, 1
        n = 0
        if last % 2 == 1:                 # odd count: the Taylor step goes alone, in place
            self.step(1, 1 % self.every == 0)
            n, self._cur, self._prev = 1, 1, 0
        self.final_index = self._cur
        return n

    def pair_advance(self, n: int, force_err: bool = False) -> int:
        """Levels n+1 and n+2 from level n; returns n + 2."""
        cur, prev = self._cur, self._prev
        out1, out2 = (2, 3) if cur in (0, 1) else (0, 1)
        l1 = step_coeffs(n + 1) + (float(self.ct[n + 1]),)
        l2 = step_coeffs(n + 2) + (float(self.ct[n + 2]),)
        w1 = force_err or (n + 1) % self.every == 0
        w2 = force_err or (n + 2) % self.every == 0
        self.plan.step_pair(n, cur, prev, out1, out2, l1, l2, self.trace[n + 1] if w1 else None,
                            self.trace[n + 2] if w2 else None)
        self.kernel_launches += 1
        if any(r >= 0 for ax in self.blk.nbr for r in ax):
            self.plan.step_fixup(n + 1, out1, cur, out2, l2, self.trace_fix[n + 2] if w2 else None)
            self.kernel_launches += 1
        self._cur, self._prev = out2, out1
        self.final_index = out2
        return n + 2

    def capture(self) -> None:
        """Capture steps 1..K into one CUDA graph (launch-bound at small blocks)."""
        if self.graph is not None:
            return
        # warm the kernels once outside capture (lazy module load is not capturable)
        self.init()
        self._enqueue_steps(1, min(2, self.p.K))
        self.barrier()
        g = torch.cuda.CUDAGraph()
        self.init()
        l0 = self.kernel_launches
        with torch.cuda.graph(g):
            self._enqueue_steps(1, self.p.K)
        self._launches_per_run = self.kernel_launches - l0
        self.kernel_launches = l0
        self.graph = g
        self.barrier()

    def run_steps(self) -> None:
        """Enqueue all K steps on the current stream (graph replay if captured)."""
        if self.use_graph and self.graph is not None:
            self.graph.replay()
            self.kernel_launches += self._launches_per_run
        else:
            self._enqueue_steps(1, self.p.K)

    def check_status(self) -> None:
        st = int(self.status[0].item())
        if st != 0:
            raise RuntimeError(f"rank {self.rank}: device watchdog tripped (status 0x{st:x}): "
                               "a neighbour's halo flag never arrived")

    def gather_trace(self) -> List[Tuple[int, float, float, float]]:
        """Combine the per-rank (max, sum sq) rows.

        Every rank's kernel has already stored its row into every other rank's table
        (P2P stores from the error epilogue), so this is a purely local read — call it
        after :meth:`barrier`.  The sum is taken in fixed rank order."""
        tc = self.table.cpu()                            # (world, K+1, slots, 2)
        mx = tc[..., 0].amax(dim=(0, 2))
        s2 = torch.zeros(tc.shape[1], dtype=torch.float64)
        for r in range(self.world):                      # fixed rank / slot order
            for q in range(tc.shape[2]):
                s2 += tc[r, :, q, 1]
        p = self.p
        return [(n, n * p.tau, float(mx[n]), math.sqrt(float(s2[n]) / p.n_unknowns))
                for n in range(1, p.K + 1) if n % self.every == 0]

    def run(self, verbose: bool = False) -> RunResult:
        """Whole run with device timing (CUDA events, max over ranks)."""
        p = self.p
        if self.use_graph and self.graph is None:
            self.capture()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        self.barrier()
        ev[0].record()
        self.plan.init(zero_second=self._zero_second)
        self.table[self.rank].zero_()
        self.barrier()          # flags must be reset everywhere before anyone pushes
        ev[1].record()
        self.run_steps()
        ev[2].record()
        self.barrier()
        self.check_status()
        t_init = ev[0].elapsed_time(ev[1]) / 1e3
        t_loop = ev[1].elapsed_time(ev[2]) / 1e3
        tt = torch.tensor([t_init, t_loop], dtype=torch.float64, device=self.device)
        if self.world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX, group=self.group)
        t_init, t_loop = (float(v) for v in tt.cpu())
        trace = self.gather_trace()
        res = RunResult(backend="fused", N=p.N, L=p.L, tau=p.tau, K=p.K, bc=p.bc, world_size=self.world,
                        decomp=self.dims, trace=trace, updates=p.updates,
                        phases={"total": t_init + t_loop, "init": t_init, "compute": t_loop, "loop": t_loop,
                                "h2d_d2h": 0.0, "exchange": 0.0},
                        extra={"plan": self.describe(), "graph": bool(self.use_graph)})
        if verbose and self.rank == 0:
            for line in res.lines(self.every):
                print(line)
        return res

    def field(self, parity: int) -> torch.Tensor:
        xo = cuda.XOFF + 1
        return self.f[parity][1:-1, 1:-1, xo: xo + self.blk.n[0]].contiguous()

    def final_field(self) -> torch.Tensor:
        """Owned unknowns of the last level computed by run_steps (any mode)."""
        if self.pair and self.final_index is not None:
            arr = (self.f + self.aux)[self.final_index]
            xo = cuda.XOFF + 1
            return arr[1:-1, 1:-1, xo: xo + self.blk.n[0]].contiguous()
        return self.field(self.p.K & 1)

    # -- tracing ---------------------------------------------------------------
    def profile_steps(self) -> List[float]:
        """Device time of every step (ms, CUDA events, eager launches; max over ranks).

        The per-step counterpart of the reference's summed phase timers (report.pdf
        p.16 §4.4); feed the result to ``utils.timing.write_chrome_trace``."""
        self.init()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(self.p.K + 1)]
        ev[0].record()
        for n in range(1, self.p.K + 1):
            self.step(n, n % self.every == 0)
            ev[n].record()
        self.barrier()
        ms = torch.tensor([ev[n - 1].elapsed_time(ev[n]) for n in range(1, self.p.K + 1)],
                          dtype=torch.float64, device=self.device)
        if self.world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX, group=self.group)
        return [float(v) for v in ms.cpu()]

    # -- checkpoint / restart ----------------------------------------------------
    def run_range(self, first: int, last: int) -> None:
        """Eagerly run steps first..last (levels first-1 -> last)."""
        self._enqueue_steps(first, last)

    def save_checkpoint(self, path: str, level: int) -> str:
        """Dump both local arrays (ghosts included) after time level ``level``."""
        from ..utils import checkpoint
        self.barrier()
        meta = {"N": self.p.N, "L": self.p.L, "tau": self.p.tau, "K": self.p.K, "bc": self.p.bc,
                "dims": self.dims, "rank": self.rank, "n": self.blk.n}
        out = checkpoint.save_state(path, self.rank, level, self.f, meta)
        self.barrier()
        return out

    def load_checkpoint(self, path: str) -> int:
        """Restore a dump; returns the level to continue from (next step = level + 1)."""
        from ..utils import checkpoint
        st = checkpoint.load_state(path, self.rank)
        m = st["meta"]
        if tuple(m["N"]) != self.p.N or tuple(m["dims"]) != self.dims or m["bc"] != self.p.bc or \
                tuple(m["n"]) != self.blk.n or abs(m["tau"] - self.p.tau) > 0:
            raise ValueError(f"checkpoint {path} does not match this problem/decomposition: {m}")
        level = int(st["level"])
        self.plan.init()                       # clears status, counters
        for dst, src in zip(self.f, st["fields"]):
            dst.copy_(src)
        # ghosts of `level` are part of the dump: tell the kernels they have arrived
        self.flags[:6].fill_(level)
        self.barrier()
        return level


class StreamingSession:
    """Step-by-step public API: the host feeds each step and reads each step's result.

    ``session.step(n)`` (1) copies that step's time factor cos(a_t t_n) from pinned
    host memory to the device table the kernel reads, (2) launches the fused step,
    (3) copies the step's (max error, sum of squares) back to pinned host memory and
    waits for it.  ``step_async(n)`` / ``result(n)`` split (3) off so that a caller can
    keep one step in flight while it consumes the previous step's result.  This is the end-to-end path a user drives when they want the error
    of every step as it happens, the way the reference prints its ``Step`` lines while
    running (report.pdf p.15-16); ``bench.py`` times it as the ``e2e`` number.
    """

    H2D_BYTES_PER_STEP = 8
    D2H_BYTES_PER_STEP = 16 * _SLOTS

    def __init__(self, solver: FusedSolver):
        self.s = solver
        p = solver.p
        self.ct_host = torch.from_numpy(solver.ct.copy()).pin_memory()
        self.ct_dev = torch.zeros(p.K + 2, dtype=torch.float64, device=solver.device)
        self.err_host = torch.zeros((p.K + 1, _SLOTS, 2), dtype=torch.float64).pin_memory()
        self._done: Dict[int, torch.cuda.Event] = {}
        solver.plan.set_ct_table(self.ct_dev)

    def close(self) -> None:
        self.s.plan.set_ct_table(None)

    def start(self) -> None:
        """Upload the analytic tables from pinned host memory and build layer 0."""
        s = self.s
        from .decomp import local_tables
        self.tab_host = [torch.from_numpy(t).pin_memory() for t in local_tables(s.p, s.blk)]
        for dst, src in zip(s.tab, self.tab_host):
            dst.copy_(src, non_blocking=True)
        s.init()

    @property
    def table_bytes(self) -> int:
        return sum(t.numel() * 8 for t in self.s.tab)

    def step_async(self, n: int) -> None:
        """Enqueue step n with its host input and the read-back of its result; do not wait.

        In pair mode two steps travel together: the call for the first step of a pair only
        uploads its input, the call for the second one launches the pass and reads both back."""
        s = self.s
        self.ct_dev[n: n + 1].copy_(self.ct_host[n: n + 1], non_blocking=True)
        if s.pair:
            if n == 1:
                self._level = s.pair_begin(s.p.K)       # runs step 1 alone when K is odd
                if self._level == 1:
                    self._finish(1, 1)
                    return
            if n == self._level + 2:
                self._level = s.pair_advance(self._level, force_err=True)
                self._finish(n - 1, n)
            return
        s.step(n, True)
        self._finish(n, n)

    def _finish(self, first: int, last: int) -> None:
        self.err_host[first: last + 1].copy_(self.s.table[self.s.rank, first: last + 1], non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        for n in range(first, last + 1):
            self._done[n] = ev

    def result(self, n: int) -> Tuple[float, float]:
        """(max error, sum of squares) of step n on this rank; waits for its read-back only."""
        if n not in self._done:
            raise RuntimeError(f"step {n} has not been launched yet (pair mode launches two steps at a time)")
        self._done.pop(n).synchronize()
        e = self.err_host[n]
        return float(e[:, 0].max()), float(e[:, 1].sum())

    def step(self, n: int) -> Tuple[float, float]:
        """Synchronous single step (not available in pair mode, where steps come in twos)."""
        if self.s.pair:
            raise RuntimeError("pair mode: use step_async(n) for consecutive n and result(n)")
        self.step_async(n)
        return self.result(n)
