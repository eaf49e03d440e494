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
  * by default two time steps share one pass over memory AND one launch AND one
    halo handshake (``pair=True``: ``w3d_step_pair`` on arrays with two ghost
    layers); ``pair=False`` runs one ``w3d_step_fused`` launch per step;
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

_FLAG_BYTES = 256     # 6 face + 12 edge flag words + padding, at the head of the symmetric buffer
_SLOTS = 2            # error-table entries per rank and level (step kernel, fix-up pass)


class _Layout:
    """Where the arrays of one rank's block sit inside its (symmetric) buffer.

    ``wide`` (pair mode): every array is allocated with two ghost layers per side in y and z
    (x as well when an x face has a neighbour) and the library works on the ``[1:-1, 1:-1]``
    view, so local index (0,0,0) and the ghost-depth-1 kernels are unchanged while the
    two-steps-per-pass kernel reaches rows / planes -1 and n+2."""

    def __init__(self, n: Sequence[int], wide: bool, x_wide: bool):
        self.n = tuple(n)
        self.wide = wide
        self.shape = cuda.wide_alloc_shape(n, x_wide) if wide else cuda.field_shape(n)
        self.pitch = self.shape[2]
        self.plane = self.shape[1] * self.shape[2]
        self.nbytes = self.shape[0] * self.plane * 8
        self.origin = (self.plane + self.pitch) * 8 if wide else 0     # byte offset of local index (0,0,0)

    def view(self, alloc: torch.Tensor) -> torch.Tensor:
        return alloc[1:-1, 1:-1] if self.wide else alloc


class FusedSolver:
    def __init__(self, p: WaveProblem, dims: Sequence[int] = (1, 1, 1), device: Optional[torch.device] = None,
                 group=None, every: int = 2, variant: int = -1, zchunk: int = 0, use_graph: bool = True,
                 spin_timeout_ms: int = 0, pair: bool = True, pair_variant: int = -1, wide_edges: bool = True):
        """``wide_edges``: on 3-D blocks (neighbours along two or three axes) also exchange the edge lines with
        the diagonal neighbours, so that two steps share one launch and one handshake there too; ``False``
        keeps the fix-up protocol (a second launch per pair) for such blocks."""
        self._launches_per_run = p.K
        self.pair = bool(pair)
        # pair mode never reads array 1 before overwriting it (the first pass skips the previous
        # level) unless the whole run is the single Taylor step
        self._zero_second = not (self.pair and p.K >= 2)
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
        x_wide = self.pair and any(b.nbr[0][s] >= 0 for b in blocks for s in range(2))
        lay = [_Layout(b.n, self.pair, x_wide) for b in blocks]
        self.lay = lay[self.rank]
        fbytes = (max(l.nbytes for l in lay) + 255) // 256 * 256
        # error table: one row per rank, (K+1) x (max, sum sq); every rank holds a full copy
        # that the other ranks' kernels fill over NVLink
        tbytes = (self.world * (p.K + 1) * _SLOTS * 16 + 255) // 256 * 256
        f0 = _FLAG_BYTES + tbytes                     # offset of the first field array
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
        nel = self.lay.shape[0] * self.lay.plane
        self.flags = buf[:_FLAG_BYTES].view(torch.int64)
        # whole allocations (all ghost layers) and the views the library works on
        self.alloc = [buf[f0 + q * fbytes: f0 + q * fbytes + nel * 8].view(torch.float64).view(self.lay.shape)
                      for q in range(narr)]
        arrs = [self.lay.view(a) for a in self.alloc]
        self.f, self.aux = arrs[:2], arrs[2:]
        # (world, K+1, slots, 2): slot 0 = the step kernels, slot 1 = the fix-up pass of a pair
        self.table = buf[_FLAG_BYTES: _FLAG_BYTES + self.world * (p.K + 1) * _SLOTS * 16].view(torch.float64).view(
            self.world, p.K + 1, _SLOTS, 2)
        gd = 2 if self.pair else 1
        self.tab_full = [torch.from_numpy(t).to(self.device) for t in local_tables(p, blk, ghosts=gd)]
        self.tab = [t[gd - 1: t.numel() - (gd - 1)] for t in self.tab_full]
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
                pb, pl = blocks[r], lay[r]
                # my low edge fills the peer's HIGH ghost (index n+1), my high edge its LOW ghost (0)
                ghost = pb.n[d] + 1 if s == 0 else 0
                base = (base_ptrs[r] if self.world > 1 else base_ptrs[0]) + f0 + pl.origin
                peers[d * 2 + s] = cuda.PeerSpec(
                    field_ptrs=(base, base + fbytes), ghost_index=ghost, pitch=pl.pitch, plane=pl.plane,
                    flag_ptr=(base_ptrs[r] if self.world > 1 else base_ptrs[0]) + 8 * (d * 2 + (1 - s)))
                peer_aux[d * 2 + s] = (base + 2 * fbytes, base + 3 * fbytes)
        self.plan = cuda.CudaPlan(blk.n, self.f, self.tab, p.coeffs(), self.flags, self.status, self.ws,
                                  peers=peers, variant=variant, zchunk=zchunk, spin_timeout_ms=spin_timeout_ms,
                                  err_table_ptr=base_ptrs[self.rank if self.world > 1 else 0] + _FLAG_BYTES,
                                  err_table_peer_ptrs=[base_ptrs[r] + _FLAG_BYTES for r in range(self.world)
                                                       if r != self.rank] if self.world > 1 else [],
                                  ghost_depth=gd)
        # device timeline: (earliest CTA start, latest CTA end) per launch, indexed by input level
        self._tl_buf = torch.zeros((p.K + 1, 2), dtype=torch.int64, device=self.device)
        self._tl_buf[:, 0] = -1
        self.plan.set_timeline(self._tl_buf)
        self.ct = time_table(p)
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.kernel_launches = 0
        # ---- time-level bookkeeping: which array holds the newest level and the one before ----
        self.level, self._cur, self._prev = 0, 0, 1
        self.final_index = None          # array index (0..3) holding the newest level after stepping in pair mode
        self.pair_mode = cuda.PAIR_LOCAL
        if self.pair:
            # 3-D blocks: the two-step stencil also reaches the ranks diagonally across the block's edges
            edges = {}
            for pi, (a, b) in enumerate(cuda.EDGE_PAIRS):
                for sa in range(2):
                    for sb in range(2):
                        ra = blk.nbr[a][sa]
                        if ra < 0 or blk.nbr[b][sb] < 0:
                            continue
                        r = blocks[ra].nbr[b][sb]          # neighbour of the neighbour = the diagonal rank
                        pb, pl = blocks[r], lay[r]
                        base = (base_ptrs[r] if self.world > 1 else base_ptrs[0])
                        o = base + f0 + pl.origin
                        edges[4 * pi + 2 * sa + sb] = dict(
                            arr=tuple(o + q * fbytes for q in range(4)),
                            ghost=(pb.n[a] + 1 if sa == 0 else 0, pb.n[b] + 1 if sb == 0 else 0),
                            pitch=pl.pitch, plane=pl.plane,
                            flag=base + 8 * (6 + 4 * pi + 2 * (1 - sa) + (1 - sb)))
            if edges and wide_edges:
                self.plan.set_edges(edges)
            self.plan.enable_pairs(self.aux[0], self.aux[1], pair_variant, zchunk, peer_aux)
            self.pair_mode = self.plan.pair_mode

    # ------------------------------------------------------------------
    def describe(self) -> str:
        return self.plan.describe()

    def barrier(self) -> None:
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            dist.barrier(self.group)

    def device_barrier(self) -> None:
        """All ranks' streams meet here WITHOUT the host: a signal-pad barrier kernel of the symmetric-memory
        handle, enqueued on the current stream (falls back to the host barrier where the handle has none)."""
        if self.world == 1:
            return
        try:
            self.symm.barrier(channel=0)
        except (AttributeError, TypeError, RuntimeError):
            self.barrier()

    def init(self, host_sync: bool = True) -> None:
        """Layer 0 everywhere (ghosts included), zero second array, flags reset.

        All ranks must have finished this before any rank's first step may push
        into their ghosts, hence the barrier — on the host by default, or (``host_sync=False``)
        a device-side barrier on the stream, so that the host can go on enqueuing steps."""
        self.plan.init(zero_second=self._zero_second)
        self.table[self.rank].zero_()                    # only my own row: peers clear theirs
        self.level, self._cur, self._prev = 0, 0, 1
        self.final_index = 0
        if host_sync:
            self.barrier()
        else:
            self.device_barrier()

    def step(self, n: int, want_err: bool = True) -> None:
        """Produce time level n (n = 1 is the Taylor step) with the one-pass-per-step kernel.

        Works on whichever two arrays hold levels n-1 and n-2 (any of the four in pair mode)."""
        a, b, cs = step_coeffs(n)
        if n == 1 and not self._zero_second:
            # init() skipped the zero fill of array 1 (pair runs never read it); this
            # single Taylor step does read it (times b = 0), so make sure it holds finite zeros
            self.f[1].zero_()
        err = self.trace[n] if want_err else None
        if self.pair:
            self.plan.step_fused_on(n - 1, self._cur, self._prev, a, b, cs, float(self.ct[n]), err)
        else:
            self.plan.step_fused(n - 1, a, b, cs, float(self.ct[n]), err)
        self.kernel_launches += 1
        self._cur, self._prev = self._prev, self._cur
        self.level = n
        self.final_index = self._cur

    def _enqueue_steps(self, first: int, last: int) -> None:
        """Levels first..last; level first-1 must be the newest one held."""
        if self.pair:
            n = first - 1
            while n + 2 <= last:
                n = self.pair_advance(n)
            if n < last:                       # odd count: the last step goes alone, in place
                self.step(last, last % self.every == 0)
            return
        for n in range(first, last + 1):
            self.step(n, n % self.every == 0)

    def pair_advance(self, n: int, force_err: bool = False, before_fixup=None) -> int:
        """Levels n+1 and n+2 from level n in ONE pass over memory; returns n + 2.

        Array roles rotate over the four arrays: (cur, prev) -> (out1, out2) = (0,1)->(2,3),
        (3,2)->(0,1), (1,0)->(2,3), ...  With neighbours (or periodic axes) the pass also does the
        halo exchange: wide-halo protocol = this one launch (3-D blocks: edge lines to the diagonal neighbours
        included); fix-up protocol (``wide_edges=False`` on 3-D blocks) = a second launch that finishes level
        n+2 next to those faces."""
        cur, prev = self._cur, self._prev
        out1, out2 = (2, 3) if cur in (0, 1) else (0, 1)
        l1 = step_coeffs(n + 1) + (float(self.ct[n + 1]),)
        l2 = step_coeffs(n + 2) + (float(self.ct[n + 2]),)
        w1 = force_err or (n + 1) % self.every == 0
        w2 = force_err or (n + 2) % self.every == 0
        self.plan.step_pair(n, cur, prev, out1, out2, l1, l2, self.trace[n + 1] if w1 else None,
                            self.trace[n + 2] if w2 else None)
        self.kernel_launches += 1
        if self.pair_mode == cuda.PAIR_FIXUP:
            if before_fixup is not None:           # test hook: e.g. a device-side delay between the two launches
                before_fixup()
            self.plan.step_fixup(n + 1, out1, cur, out2, l2, self.trace_fix[n + 2] if w2 else None)
            self.kernel_launches += 1
        self._cur, self._prev = out2, out1
        self.level = n + 2
        self.final_index = out2
        return n + 2

    def capture(self) -> None:
        """Capture steps 1..K into one CUDA graph (launch-bound at small blocks)."""
        if self.graph is not None:
            return
        # warm the kernels once outside capture (lazy module load is not capturable)
        self.init()
        self._enqueue_steps(1, min(3, self.p.K))
        self.barrier()
        g = torch.cuda.CUDAGraph()
        self.init()
        l0 = self.kernel_launches
        with torch.cuda.graph(g):
            self._enqueue_steps(1, self.p.K)
        self._graph_state = (self.level, self._cur, self._prev, self.final_index)
        self._launches_per_run = self.kernel_launches - l0
        self.kernel_launches = l0
        self.graph = g
        self.barrier()

    def run_steps(self) -> None:
        """Enqueue all K steps on the current stream (graph replay if captured)."""
        if self.use_graph and self.graph is not None:
            self.graph.replay()
            self.kernel_launches += self._launches_per_run
            self.level, self._cur, self._prev, self.final_index = self._graph_state
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

    def exchange_stats(self) -> Dict[str, float]:
        """Exchange time since the last init(), measured on the device and max-reduced over ranks.

        ``flag_wait_s``: per launch, the longest any face CTA spun on its neighbours' flags;
        ``halo_copy_s``: per launch, the longest halo copy-out (P2P stores + system fence); both
        summed over the launches.  These are the counterparts of the reference's "MPI" and
        "H2D/D2H" columns (report.pdf p.16 §4.4) — here they run underneath the interior CTAs."""
        st = self.plan.exchange_stats()
        if self.world > 1:
            t = torch.tensor([st["flag_wait_s"], st["halo_copy_s"], st["max_flag_wait_s"], st["max_halo_copy_s"]],
                             dtype=torch.float64, device=self.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
            st["flag_wait_s"], st["halo_copy_s"], st["max_flag_wait_s"], st["max_halo_copy_s"] = (float(v) for v in t.cpu())
        return st

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
        self.level, self._cur, self._prev, self.final_index = 0, 0, 1, 0
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
        xs = self.exchange_stats()
        trace = self.gather_trace()
        # `exchange` (flag waits + halo copy-out of the face CTAs) is measured inside the kernels and
        # runs concurrently with the other CTAs: it is contained in `loop`, not added to it
        res = RunResult(backend="fused", N=p.N, L=p.L, tau=p.tau, K=p.K, bc=p.bc, world_size=self.world,
                        decomp=self.dims, trace=trace, updates=p.updates,
                        phases={"total": t_init + t_loop, "init": t_init, "compute": t_loop, "loop": t_loop,
                                "h2d_d2h": 0.0, "exchange": xs["flag_wait_s"] + xs["halo_copy_s"],
                                "flag_wait": xs["flag_wait_s"], "halo_copy": xs["halo_copy_s"]},
                        extra={"plan": self.describe(), "graph": bool(self.use_graph), "exchange_stats": xs})
        if verbose and self.rank == 0:
            for line in res.lines(self.every):
                print(line)
        return res

    def _owned(self, arr: torch.Tensor) -> torch.Tensor:
        xo = cuda.XOFF + 1
        return arr[1:-1, 1:-1, xo: xo + self.blk.n[0]].contiguous()

    def field(self, parity: int) -> torch.Tensor:
        return self._owned(self.f[parity])

    def final_field(self) -> torch.Tensor:
        """Owned unknowns of the newest level computed so far (any mode)."""
        if self.pair and self.final_index is not None:
            return self._owned((self.f + self.aux)[self.final_index])
        return self.field(self.level & 1)

    # -- tracing ---------------------------------------------------------------
    def device_timeline(self) -> List[Tuple[int, float, float]]:
        """One whole run with %globaltimer stamps taken inside the kernels.

        Returns [(input level, start_us, end_us)] per launch on this rank, relative to the first
        start: start = earliest CTA of the launch, end = latest CTA.  end(p) -> start(p+1) is the
        launch gap; what lies between start and end includes the flag waits (see exchange_stats).
        Uses the captured CUDA graph if there is one (the stamp buffer is part of every launch)."""
        self.init()
        self._tl_buf[:, 0] = -1           # 0xffff... = UINT64_MAX for the device-side atomicMin
        self._tl_buf[:, 1] = 0
        self.run_steps()
        self.barrier()
        h = self._tl_buf.cpu()
        rows = [(s, int(h[s, 0]), int(h[s, 1])) for s in range(h.shape[0]) if int(h[s, 1]) != 0]
        t0 = min(r[1] for r in rows)
        return [(s, (a - t0) / 1e3, (b - t0) / 1e3) for s, a, b in rows]

    def profile_steps(self) -> List[float]:
        """Device time per step (ms, CUDA events, eager launches; max over ranks).

        In pair mode two steps share a pass: each of them is reported with half the pass time.
        The per-step counterpart of the reference's summed phase timers (report.pdf
        p.16 §4.4); feed the result to ``utils.timing.write_chrome_trace``."""
        self.init()
        K = self.p.K
        marks = [(0, torch.cuda.Event(enable_timing=True))]
        marks[0][1].record()
        n = 0
        while n < K:
            if self.pair and n + 2 <= K:
                n = self.pair_advance(n)
            else:
                n += 1
                self.step(n, n % self.every == 0)
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            marks.append((n, e))
        self.barrier()
        ms = []
        for (n0, e0), (n1, e1) in zip(marks[:-1], marks[1:]):
            ms += [e0.elapsed_time(e1) / (n1 - n0)] * (n1 - n0)
        t = torch.tensor(ms, dtype=torch.float64, device=self.device)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return [float(v) for v in t.cpu()]

    # -- checkpoint / restart ----------------------------------------------------
    def run_range(self, first: int, last: int) -> None:
        """Eagerly run steps first..last (levels first-1 -> last)."""
        if first - 1 != self.level:
            raise ValueError(f"run_range({first}, {last}): the solver holds level {self.level}")
        self._enqueue_steps(first, last)

    def save_checkpoint(self, path: str, level: Optional[int] = None) -> str:
        """Dump the two live time levels (all ghost layers included) after the newest level.

        In pair mode the live levels may sit in any two of the four arrays; the dump records
        which, and a restart puts them back into the same arrays."""
        from ..utils import checkpoint
        self.barrier()
        if level is not None and level != self.level:
            raise ValueError(f"checkpoint asked for level {level} but the solver holds level {self.level}")
        meta = {"N": self.p.N, "L": self.p.L, "tau": self.p.tau, "K": self.p.K, "bc": self.p.bc,
                "dims": self.dims, "rank": self.rank, "n": self.blk.n, "pair": self.pair,
                "roles": (self._cur, self._prev), "shape": tuple(self.lay.shape)}
        out = checkpoint.save_state(path, self.rank, self.level, [self.alloc[self._cur], self.alloc[self._prev]], meta)
        self.barrier()
        return out

    def load_checkpoint(self, path: str) -> int:
        """Restore a dump; returns the level to continue from (next step = level + 1)."""
        from ..utils import checkpoint
        st = checkpoint.load_state(path, self.rank)
        m = st["meta"]
        if tuple(m["N"]) != self.p.N or tuple(m["dims"]) != self.dims or m["bc"] != self.p.bc or \
                tuple(m["n"]) != self.blk.n or abs(m["tau"] - self.p.tau) > 0 or \
                bool(m.get("pair", False)) != self.pair or tuple(m.get("shape", self.lay.shape)) != tuple(self.lay.shape):
            raise ValueError(f"checkpoint {path} does not match this problem/decomposition/mode: {m}")
        level = int(st["level"])
        cur, prev = m.get("roles", (level & 1, (level & 1) ^ 1))
        self.plan.init()                       # clears status, counters
        self.alloc[cur].copy_(st["fields"][0])
        self.alloc[prev].copy_(st["fields"][1])
        self.level, self._cur, self._prev, self.final_index = level, cur, prev, cur
        # ghosts of `level` are part of the dump: tell the kernels they have arrived
        self.flags[:cuda.FLAG_WORDS].fill_(level)
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
        self.tab_host = [torch.from_numpy(t).pin_memory()
                         for t in local_tables(p, solver.blk, ghosts=2 if solver.pair else 1)]
        # The per-step host work (8-byte upload, choice of arrays, launch, read-back, event) runs in native
        # code (csrc/cuda/session.cu): a pass is 105 us at 512^3 on eight GPUs, an interpreter is slower.
        self.native = cuda.NativeSession(solver.plan, p.K, _SLOTS, self.ct_host, self.ct_dev,
                                         solver.table[solver.rank], self.err_host)

    def close(self) -> None:
        """Detach from the plan; the solver takes over where the session stopped."""
        if self.native is not None:
            self._sync_solver()
            self.native.close()
            self.native = None

    def _sync_solver(self) -> None:
        s = self.s
        s.level, s._cur, s._prev = self.native.stepper()
        s.final_index = s._cur

    def start(self) -> None:
        """Upload the analytic tables from pinned host memory and build layer 0."""
        s = self.s
        for dst, src in zip(s.tab_full, self.tab_host):
            dst.copy_(src, non_blocking=True)
        s.init(host_sync=False)
        self.native.begin(0, 0, 1)
        self._stream = cuda.stream_ptr()

    @property
    def table_bytes(self) -> int:
        return sum(t.numel() * 8 for t in self.s.tab_full)

    def step_async(self, n: int) -> None:
        """Enqueue step n with its host input and the read-back of its result; do not wait.

        In pair mode two steps travel together: the call for the first step of a pair only
        uploads its input, the call for the second one launches the pass and reads both back;
        the last step of an odd-length run goes alone."""
        self.native.step_async(n, self._stream)

    def result(self, n: int) -> Tuple[float, float]:
        """(max error, sum of squares) of step n on this rank; waits for its read-back only."""
        return self.native.result(n)

    def step(self, n: int) -> Tuple[float, float]:
        """Synchronous single step (not available in pair mode, where steps come in twos)."""
        if self.s.pair:
            raise RuntimeError("pair mode: use step_async(n) for consecutive n and result(n)")
        self.step_async(n)
        return self.result(n)
