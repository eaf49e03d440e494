"""ctypes binding of libwave3d_b200.so (include/wave3d.h) — the sm_100a kernels.

The library is plain C ABI built in-tree by nvcc (no torch headers): Python hands
it raw device pointers of torch tensors and the current CUDA stream handle, so
every launch lands on torch's stream and can be captured in a CUDA graph.

Reference parity: these are the kernels the reference's mpigpu-1.cu needed
(report.pdf p.14-17 §4), re-designed for B200.
There is deliberately NO eager/PyTorch fallback here: if the extension cannot be
loaded on a GPU box the import fails loudly.
"""
from __future__ import annotations

import ctypes as C
from functools import lru_cache
from typing import Optional, Sequence, Tuple

import torch

from ..build import CUDA_LIB, ensure
from .cpu import W3dBlock

XOFF = 15
PITCH_ALIGN = 16
ERR_SLOTS = 8


class W3dPeer(C.Structure):
    _fields_ = [("active", C.c_int32), ("ghost_index", C.c_int32), ("field", C.c_void_p * 2),
                ("pitch", C.c_int64), ("plane", C.c_int64), ("flag", C.c_void_p), ("err_slot", C.c_void_p)]


class W3dPlanDesc(C.Structure):
    _fields_ = [("blk", W3dBlock), ("field", C.c_void_p * 2), ("tab", C.c_void_p * 3),
                ("coef", C.c_double * 3), ("peer", W3dPeer * 6), ("flags_in", C.c_void_p),
                ("status", C.c_void_p), ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t),
                ("variant", C.c_int32), ("zchunk", C.c_int32), ("num_sms", C.c_int32),
                ("spin_timeout_ms", C.c_int32), ("err_table", C.c_void_p), ("err_table_peer", C.c_void_p * 7),
                ("n_err_peers", C.c_int32), ("ghost_depth", C.c_int32), ("plane_stride", C.c_int64)]


class W3dEdgePeer(C.Structure):
    """The rank diagonally across one edge of the block (include/wave3d.h: w3d_plan_set_edges)."""
    _fields_ = [("active", C.c_int32), ("ghost_a", C.c_int32), ("ghost_b", C.c_int32), ("arr", C.c_void_p * 4),
                ("pitch", C.c_int64), ("plane", C.c_int64), ("flag", C.c_void_p)]


EDGE_PAIRS = ((0, 1), (0, 2), (1, 2))      # edge e = 4 * pair + 2 * sa + sb
FLAG_WORDS = 18                            # 6 faces + 12 edges


class W3dStepper(C.Structure):
    """Which arrays hold the newest level and the one before (include/wave3d.h)."""
    _fields_ = [("level", C.c_int32), ("cur", C.c_int32), ("prev", C.c_int32), ("reserved", C.c_int32)]


class W3dSessionDesc(C.Structure):
    _fields_ = [("plan", C.c_void_p), ("K", C.c_int32), ("slots", C.c_int32), ("ct_host", C.c_void_p),
                ("ct_dev", C.c_void_p), ("err_dev", C.c_void_p), ("err_host", C.c_void_p)]


class CudaLibError(RuntimeError):
    pass


@lru_cache(maxsize=1)
def lib() -> C.CDLL:
    path = ensure(CUDA_LIB)
    l = C.CDLL(str(path))
    vp, i32p, dbl = C.c_void_p, C.POINTER(C.c_int32), C.c_double
    l.w3d_last_error.restype = C.c_char_p
    l.w3d_version.restype = C.c_char_p
    l.w3d_workspace_bytes.restype = C.c_size_t
    l.w3d_error_workspace_bytes.restype = C.c_size_t
    l.w3d_pitch_for.restype = C.c_int64
    l.w3d_pitch_for.argtypes = [C.c_int32]
    l.w3d_plan_create.argtypes = [C.POINTER(W3dPlanDesc), C.POINTER(vp)]
    l.w3d_plan_destroy.argtypes = [vp]
    l.w3d_plan_destroy.restype = None
    l.w3d_plan_describe.argtypes = [vp]
    l.w3d_plan_describe.restype = C.c_char_p
    l.w3d_plan_launches.argtypes = [vp]
    l.w3d_plan_launches.restype = C.c_longlong
    l.w3d_plan_init.argtypes = [vp, vp]
    l.w3d_plan_init_ex.argtypes = [vp, vp, C.c_int]
    l.w3d_plan_set_ct_table.argtypes = [vp, vp]
    l.w3d_step_fused.argtypes = [vp, C.c_int, dbl, dbl, dbl, dbl, vp, vp]
    l.w3d_step_fused_on.argtypes = [vp, C.c_int, C.c_int, C.c_int, dbl, dbl, dbl, dbl, vp, vp]
    l.w3d_plan_pair_mode.argtypes = [vp]
    l.w3d_plan_pairs_enabled.argtypes = [vp]
    l.w3d_plan_set_edges.argtypes = [vp, C.POINTER(W3dEdgePeer)]
    l.w3d_stepper_reset.argtypes = [C.POINTER(W3dStepper)]
    l.w3d_stepper_reset.restype = None
    l.w3d_advance_pair.argtypes = [vp, C.POINTER(W3dStepper), vp, vp, vp, vp, vp]
    l.w3d_advance_single.argtypes = [vp, C.POINTER(W3dStepper), vp, vp, vp]
    l.w3d_session_create.argtypes = [C.POINTER(W3dSessionDesc), C.POINTER(vp)]
    l.w3d_session_destroy.argtypes = [vp]
    l.w3d_session_destroy.restype = None
    l.w3d_session_begin.argtypes = [vp, C.POINTER(W3dStepper)]
    l.w3d_session_step_async.argtypes = [vp, C.c_int, vp]
    l.w3d_session_result.argtypes = [vp, C.c_int, C.POINTER(C.c_double)]
    l.w3d_session_stepper.argtypes = [vp, C.POINTER(W3dStepper)]
    l.w3d_plan_set_timeline.argtypes = [vp, vp, C.c_int]
    l.w3d_plan_exchange_stats.argtypes = [vp, C.POINTER(C.c_ulonglong), vp]
    l.w3d_step_tma.argtypes = [vp, C.c_int, dbl, dbl, dbl, dbl, i32p, i32p, vp, C.c_int, vp]
    l.w3d_step_box.argtypes = [vp, C.c_int, dbl, dbl, dbl, dbl, i32p, i32p, vp, C.c_int, vp]
    l.w3d_plan_enable_pairs.argtypes = [vp, vp, vp, C.POINTER(vp), C.c_int, C.c_int]
    l.w3d_step_pair.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, dbl, dbl, dbl, dbl, dbl, dbl, dbl, dbl, vp, vp, vp]
    l.w3d_step_fixup.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, dbl, dbl, dbl, dbl, vp, vp]
    l.w3d_pack_face.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp]
    l.w3d_unpack_face.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp]
    l.w3d_wrap_axis.argtypes = [vp, C.c_int, C.c_int, vp]
    l.w3d_error_norms.argtypes = [vp, C.POINTER(W3dBlock), vp, vp, vp, dbl, vp, vp, C.c_size_t, vp]
    return l


def check(rc: int) -> None:
    if rc != 0:
        raise CudaLibError(f"libwave3d_b200 error {rc}: {lib().w3d_last_error().decode()}")


PAIR_LOCAL, PAIR_WIDE, PAIR_FIXUP = 0, 1, 2      # w3d_plan_pair_mode()


def pitch_for(nx: int, x_ghosts: int = 1) -> int:
    """Row pitch (elements) for nx unknowns and `x_ghosts` ghost columns per side."""
    need = XOFF + nx + 1 + x_ghosts
    return (need + PITCH_ALIGN - 1) // PITCH_ALIGN * PITCH_ALIGN


def make_block(n: Sequence[int], pitch: int = 0) -> W3dBlock:
    b = W3dBlock()
    b.n[:] = list(n)
    b.xoff = XOFF
    b.pitch = pitch or pitch_for(n[0])
    return b


def field_shape(n: Sequence[int]) -> Tuple[int, int, int]:
    return (n[2] + 2, n[1] + 2, pitch_for(n[0]))


def wide_alloc_shape(n: Sequence[int], x_wide: bool = False) -> Tuple[int, int, int]:
    """Allocation with TWO ghost layers per side in y and z (and x when ``x_wide``).

    The library is handed the view ``alloc[1:-1, 1:-1]`` — local index (0,0,0) as in the classic
    layout, explicit plane stride — so every kernel written for one ghost layer runs on it
    unchanged, and the two-steps-per-pass kernel reaches rows / planes -1 and n+2."""
    return (n[2] + 4, n[1] + 4, pitch_for(n[0], 2 if x_wide else 1))


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def workspace_bytes() -> int:
    return int(lib().w3d_workspace_bytes())


def _box(v: Optional[Sequence[int]]):
    return None if v is None else (C.c_int32 * 3)(*v)


class PeerSpec:
    """Where one of my faces has to be delivered (all integers are device addresses)."""

    def __init__(self, field_ptrs: Sequence[int], ghost_index: int, pitch: int, plane: int, flag_ptr: int):
        self.field_ptrs = tuple(field_ptrs)
        self.ghost_index = ghost_index
        self.pitch = pitch
        self.plane = plane
        self.flag_ptr = flag_ptr


class CudaPlan:
    """Owns a W3dPlan handle plus the torch tensors it points into."""

    def __init__(self, n: Sequence[int], fields: Sequence[torch.Tensor], tables: Sequence[torch.Tensor],
                 coef: Sequence[float], flags_in: torch.Tensor, status: torch.Tensor,
                 workspace: torch.Tensor, peers: Optional[dict] = None, variant: int = -1,
                 zchunk: int = 0, spin_timeout_ms: int = 0, err_table_ptr: int = 0,
                 err_table_peer_ptrs: Sequence[int] = (), ghost_depth: int = 1):
        """fields: two tensors of shape (nz+2, ny+2, pitch).  With ``ghost_depth=2`` they are the
        ``[1:-1, 1:-1]`` views of allocations made with :func:`wide_alloc_shape`, and the tables
        are the ``[1:-1]`` views of tables with n+4 entries."""
        self.n = tuple(int(x) for x in n)
        self.fields = list(fields)
        self.tables = list(tables)
        self.flags_in, self.status, self.workspace = flags_in, status, workspace
        self.ghost_depth = ghost_depth
        f0 = self.fields[0]
        pitch, plane = int(f0.stride(1)), int(f0.stride(0))
        self.blk = make_block(self.n, pitch)
        for f in self.fields:
            assert f.is_cuda and f.dtype == torch.float64 and tuple(f.shape[:2]) == (self.n[2] + 2, self.n[1] + 2)
            assert f.stride(2) == 1 and f.stride(1) == pitch and f.stride(0) == plane and f.shape[2] == pitch
            if ghost_depth < 2:
                assert f.is_contiguous()
        for t, nn in zip(self.tables, self.n):
            assert t.is_cuda and t.dtype == torch.float64 and t.numel() == nn + 2
        d = W3dPlanDesc()
        d.blk = self.blk
        d.ghost_depth = ghost_depth
        d.plane_stride = plane
        d.field[0], d.field[1] = self.fields[0].data_ptr(), self.fields[1].data_ptr()
        for k in range(3):
            d.tab[k] = self.tables[k].data_ptr()
            d.coef[k] = float(coef[k])
        for f, spec in (peers or {}).items():
            pr = d.peer[f]
            pr.active = 1
            pr.ghost_index = spec.ghost_index
            pr.field[0], pr.field[1] = spec.field_ptrs
            pr.pitch, pr.plane = spec.pitch, spec.plane
            pr.flag = spec.flag_ptr
        d.flags_in = flags_in.data_ptr()
        d.status = status.data_ptr()
        d.workspace = workspace.data_ptr()
        d.workspace_bytes = workspace.numel() * workspace.element_size()
        d.variant, d.zchunk, d.num_sms, d.spin_timeout_ms = variant, zchunk, 0, spin_timeout_ms
        if err_table_ptr and err_table_peer_ptrs:
            if len(err_table_peer_ptrs) > 7:
                raise ValueError("at most 7 peer error tables (8 GPUs per node)")
            d.err_table = err_table_ptr
            for q, ptr in enumerate(err_table_peer_ptrs):
                d.err_table_peer[q] = ptr
            d.n_err_peers = len(err_table_peer_ptrs)
        self._desc = d
        # plan creation copies a small table into the workspace with a blocking copy: make sure
        # the fills that created the tensors (possibly on another stream) are done first
        torch.cuda.synchronize()
        h = C.c_void_p()
        check(lib().w3d_plan_create(C.byref(d), C.byref(h)))
        self._h = h

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            try:
                lib().w3d_plan_destroy(h)
            except Exception:          # interpreter shutdown: the module globals may already be gone
                pass
            self._h = None

    def describe(self) -> str:
        return lib().w3d_plan_describe(self._h).decode()

    @property
    def launches(self) -> int:
        return int(lib().w3d_plan_launches(self._h))

    def init(self, zero_second: bool = True) -> None:
        check(lib().w3d_plan_init_ex(self._h, stream_ptr(), int(zero_second)))

    def set_ct_table(self, table: Optional[torch.Tensor]) -> None:
        self._ct_table = table
        check(lib().w3d_plan_set_ct_table(self._h, None if table is None else table.data_ptr()))

    @staticmethod
    def _err_ptr(err_out: Optional[torch.Tensor]):
        return None if err_out is None else err_out.data_ptr()

    def step_fused(self, s: int, a: float, b: float, cscale: float, ct: float,
                   err_out: Optional[torch.Tensor] = None) -> None:
        check(lib().w3d_step_fused(self._h, s, a, b, cscale, ct, self._err_ptr(err_out), stream_ptr()))

    def step_tma(self, s, a, b, cscale, ct, lo=None, hi=None, err_out=None, slot=0) -> None:
        check(lib().w3d_step_tma(self._h, s, a, b, cscale, ct, _box(lo), _box(hi), self._err_ptr(err_out),
                                 slot, stream_ptr()))

    def step_box(self, s, a, b, cscale, ct, lo=None, hi=None, err_out=None, slot=0) -> None:
        check(lib().w3d_step_box(self._h, s, a, b, cscale, ct, _box(lo), _box(hi), self._err_ptr(err_out),
                                 slot, stream_ptr()))

    def set_edges(self, edges: dict) -> None:
        """edges: edge index -> dict(arr=(4 addresses), ghost=(ga, gb), pitch=, plane=, flag=address); before enable_pairs."""
        tab = (W3dEdgePeer * 12)()
        for e, spec in edges.items():
            q = tab[e]
            q.active = 1
            q.ghost_a, q.ghost_b = spec["ghost"]
            for i in range(4):
                q.arr[i] = spec["arr"][i]
            q.pitch, q.plane, q.flag = spec["pitch"], spec["plane"], spec["flag"]
        check(lib().w3d_plan_set_edges(self._h, tab))

    def enable_pairs(self, aux0: torch.Tensor, aux1: torch.Tensor, pair_variant: int = -1, zchunk: int = 0,
                     peer_aux: Optional[dict] = None) -> None:
        """Register two more arrays (same layout) so that two steps can be fused per pass.

        peer_aux: face index -> (aux0 address, aux1 address) of that neighbour's arrays."""
        for a in (aux0, aux1):
            assert a.is_cuda and a.dtype == torch.float64 and a.shape == self.fields[0].shape
            assert a.stride() == self.fields[0].stride()
        self.aux = [aux0, aux1]
        table = (C.c_void_p * 12)()
        for f, (p0, p1) in (peer_aux or {}).items():
            table[2 * f], table[2 * f + 1] = p0, p1
        check(lib().w3d_plan_enable_pairs(self._h, aux0.data_ptr(), aux1.data_ptr(), table, pair_variant, zchunk))

    def set_timeline(self, buf: Optional[torch.Tensor]) -> None:
        """buf: int64 tensor of shape (capacity, 2) pre-filled with (INT64 max as uint64 max, 0), or None."""
        self._timeline = buf
        check(lib().w3d_plan_set_timeline(self._h, None if buf is None else buf.data_ptr(),
                                          0 if buf is None else buf.shape[0]))

    @property
    def pair_mode(self) -> int:
        """PAIR_LOCAL / PAIR_WIDE / PAIR_FIXUP: how pair passes treat faces with a neighbour."""
        return int(lib().w3d_plan_pair_mode(self._h))

    def step_fused_on(self, s: int, cur: int, prev: int, a: float, b: float, cscale: float, ct: float,
                      err_out: Optional[torch.Tensor] = None) -> None:
        """Single step with level s in array `cur` and level s-1 in array `prev` (overwritten)."""
        check(lib().w3d_step_fused_on(self._h, s, cur, prev, a, b, cscale, ct, self._err_ptr(err_out), stream_ptr()))

    def exchange_stats(self) -> dict:
        """Device-measured exchange time since init (see w3d_plan_exchange_stats); synchronises."""
        out = (C.c_ulonglong * 5)()
        check(lib().w3d_plan_exchange_stats(self._h, out, stream_ptr()))
        return {"flag_wait_s": out[0] * 1e-9, "halo_copy_s": out[1] * 1e-9, "launches": int(out[2]),
                "max_flag_wait_s": out[3] * 1e-9, "max_halo_copy_s": out[4] * 1e-9}

    def step_pair(self, s: int, cur: int, prev: int, out1: int, out2: int, lvl1, lvl2, err1=None, err2=None) -> None:
        """s = level held by array `cur`; lvl1 / lvl2 = (a, b, cscale, ct) of the two produced levels."""
        check(lib().w3d_step_pair(self._h, s, cur, prev, out1, out2, *lvl1, *lvl2, self._err_ptr(err1),
                                  self._err_ptr(err2), stream_ptr()))

    def step_fixup(self, level: int, cur: int, prev: int, out: int, lvl, err=None) -> None:
        """Finish level+1 on the layer next to faces with a neighbour (see w3d_step_fixup)."""
        check(lib().w3d_step_fixup(self._h, level, cur, prev, out, *lvl, self._err_ptr(err), stream_ptr()))

    def pack_face(self, parity: int, axis: int, side: int, buf: torch.Tensor, ghost: bool = False) -> None:
        check(lib().w3d_pack_face(self._h, parity, axis, side, int(ghost), buf.data_ptr(), stream_ptr()))

    def unpack_face(self, parity: int, axis: int, side: int, buf: torch.Tensor, ghost: bool = True) -> None:
        check(lib().w3d_unpack_face(self._h, parity, axis, side, int(ghost), buf.data_ptr(), stream_ptr()))

    def wrap_axis(self, parity: int, axis: int) -> None:
        check(lib().w3d_wrap_axis(self._h, parity, axis, stream_ptr()))


def error_norms(u: torch.Tensor, n: Sequence[int], tables: Sequence[torch.Tensor], ct: float,
                out: torch.Tensor, workspace: torch.Tensor) -> None:
    """Stand-alone error reduction (K6): out[0] = max error, out[1] = sum of squares."""
    blk = make_block(n)
    check(lib().w3d_error_norms(u.data_ptr(), C.byref(blk), tables[0].data_ptr(), tables[1].data_ptr(),
                                tables[2].data_ptr(), ct, out.data_ptr(), workspace.data_ptr(),
                                workspace.numel() * workspace.element_size(), stream_ptr()))


class NativeSession:
    """Thin handle on a W3dSession (csrc/cuda/session.cu): the per-step host work in native code."""

    def __init__(self, plan: CudaPlan, K: int, slots: int, ct_host: torch.Tensor, ct_dev: torch.Tensor,
                 err_dev: torch.Tensor, err_host: torch.Tensor):
        assert ct_host.is_pinned() and err_host.is_pinned() and ct_dev.is_cuda and err_dev.is_cuda
        assert err_dev.is_contiguous() and err_host.is_contiguous() and err_dev.numel() == err_host.numel()
        self._keep = (plan, ct_host, ct_dev, err_dev, err_host)      # the session only borrows these
        d = W3dSessionDesc()
        d.plan, d.K, d.slots = plan._h, K, slots
        d.ct_host, d.ct_dev = ct_host.data_ptr(), ct_dev.data_ptr()
        d.err_dev, d.err_host = err_dev.data_ptr(), err_host.data_ptr()
        h = C.c_void_p()
        check(lib().w3d_session_create(C.byref(d), C.byref(h)))
        self._h = h
        self._out = (C.c_double * 2)()
        self._step, self._result = lib().w3d_session_step_async, lib().w3d_session_result

    def close(self) -> None:
        if getattr(self, "_h", None):
            lib().w3d_session_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:              # interpreter shutdown
            pass

    def begin(self, level: int = 0, cur: int = 0, prev: int = 1) -> None:
        st = W3dStepper(level, cur, prev, 0)
        check(lib().w3d_session_begin(self._h, C.byref(st)))

    def step_async(self, n: int, stream: int) -> None:
        rc = self._step(self._h, n, stream)
        if rc != 0:
            raise CudaLibError(f"w3d_session_step_async({n}) failed ({rc}): feed consecutive steps 1..K; "
                               f"{lib().w3d_last_error().decode()}")

    def result(self, n: int) -> Tuple[float, float]:
        rc = self._result(self._h, n, self._out)
        if rc != 0:
            raise CudaLibError(f"w3d_session_result({n}) failed ({rc}): step not launched yet "
                               "(pair plans launch two steps at a time)")
        return self._out[0], self._out[1]

    def stepper(self) -> Tuple[int, int, int]:
        st = W3dStepper()
        check(lib().w3d_session_stepper(self._h, C.byref(st)))
        return st.level, st.cur, st.prev
