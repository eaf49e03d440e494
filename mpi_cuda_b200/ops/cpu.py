"""ctypes binding of libwave3d_cpu.so (include/wave3d_cpu.h).

Python face of the native CPU solvers that stand in for the reference's
wave.cpp / openmpwave.cpp / mpi.cpp / mpiomp.cpp (report.pdf p.20-26 §5).
"""
from __future__ import annotations

import ctypes as C
from functools import lru_cache
from typing import List, Sequence, Tuple

import numpy as np

from ..build import CPU_LIB, ensure
from ..models.problem import WaveProblem


class W3dBlock(C.Structure):
    _fields_ = [("n", C.c_int32 * 3), ("xoff", C.c_int32), ("pitch", C.c_int64)]


class W3dCpuRun(C.Structure):
    _fields_ = [("N", C.c_int32 * 3), ("L", C.c_double * 3), ("tau", C.c_double), ("K", C.c_int32),
                ("periodic", C.c_int32 * 3), ("threads", C.c_int32), ("dims", C.c_int32 * 3),
                ("every", C.c_int32), ("verbose", C.c_int32)]


_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)


@lru_cache(maxsize=1)
def lib() -> C.CDLL:
    l = C.CDLL(str(ensure(CPU_LIB)))
    l.w3d_block_elems.restype = C.c_int64
    l.w3d_block_elems.argtypes = [C.POINTER(W3dBlock)]
    l.w3d_cpu_init.argtypes = [_dp, C.POINTER(W3dBlock), _dp, _dp, _dp, C.c_int]
    l.w3d_cpu_step.argtypes = [_dp, _dp, C.POINTER(W3dBlock), C.c_double, C.c_double, _dp, _ip, _ip, C.c_int]
    l.w3d_cpu_error.argtypes = [_dp, C.POINTER(W3dBlock), _dp, _dp, _dp, C.c_double, _dp, _dp, C.c_int]
    l.w3d_cpu_wrap.argtypes = [_dp, C.POINTER(W3dBlock), C.c_int]
    l.w3d_cpu_face_elems.restype = C.c_int64
    l.w3d_cpu_face_elems.argtypes = [C.POINTER(W3dBlock), C.c_int]
    l.w3d_cpu_pack.argtypes = [_dp, C.POINTER(W3dBlock), C.c_int, C.c_int, C.c_int, _dp]
    l.w3d_cpu_unpack.argtypes = [_dp, C.POINTER(W3dBlock), C.c_int, C.c_int, C.c_int, _dp]
    l.w3d_cpu_solve.restype = C.c_int
    l.w3d_cpu_solve.argtypes = [C.POINTER(W3dCpuRun), _dp, _dp]
    l.w3d_cpu_version.restype = C.c_char_p
    l.w3d_cpu_max_threads.restype = C.c_int
    return l


def _p(a: np.ndarray):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_dp)


def make_block(n: Sequence[int], xoff: int = 0, pitch: int | None = None) -> W3dBlock:
    b = W3dBlock()
    b.n[:] = list(n)
    b.xoff = xoff
    b.pitch = pitch if pitch is not None else xoff + n[0] + 2
    return b


def block_shape(b: W3dBlock) -> Tuple[int, int, int]:
    return (b.n[2] + 2, b.n[1] + 2, int(b.pitch))


def init(u: np.ndarray, b: W3dBlock, sx, sy, sz, threads: int = 1) -> None:
    lib().w3d_cpu_init(_p(u), C.byref(b), _p(sx), _p(sy), _p(sz), threads)


def step(prev_next: np.ndarray, cur: np.ndarray, b: W3dBlock, a: float, bb: float, c: Sequence[float],
         lo: Sequence[int] | None = None, hi: Sequence[int] | None = None, threads: int = 1) -> None:
    lo = [1, 1, 1] if lo is None else list(lo)
    hi = [b.n[0] + 1, b.n[1] + 1, b.n[2] + 1] if hi is None else list(hi)
    lib().w3d_cpu_step(_p(prev_next), _p(cur), C.byref(b), a, bb, (C.c_double * 3)(*c),
                       (C.c_int32 * 3)(*lo), (C.c_int32 * 3)(*hi), threads)


def error(u: np.ndarray, b: W3dBlock, sx, sy, sz, ct: float, threads: int = 1) -> Tuple[float, float]:
    mx, s2 = C.c_double(), C.c_double()
    lib().w3d_cpu_error(_p(u), C.byref(b), _p(sx), _p(sy), _p(sz), ct, C.byref(mx), C.byref(s2), threads)
    return mx.value, s2.value


def wrap(u: np.ndarray, b: W3dBlock, axis: int) -> None:
    lib().w3d_cpu_wrap(_p(u), C.byref(b), axis)


def face_elems(b: W3dBlock, axis: int) -> int:
    return int(lib().w3d_cpu_face_elems(C.byref(b), axis))


def pack(u: np.ndarray, b: W3dBlock, axis: int, side: int, ghost: bool = False) -> np.ndarray:
    buf = np.empty(face_elems(b, axis), dtype=np.float64)
    lib().w3d_cpu_pack(_p(u), C.byref(b), axis, side, int(ghost), _p(buf))
    return buf


def unpack(u: np.ndarray, b: W3dBlock, axis: int, side: int, buf: np.ndarray, ghost: bool = True) -> None:
    lib().w3d_cpu_unpack(_p(u), C.byref(b), axis, side, int(ghost), _p(np.ascontiguousarray(buf)))


def solve(p: WaveProblem, threads: int = 1, dims: Sequence[int] = (1, 1, 1), every: int = 1,
          verbose: bool = False):
    """Whole run in native code. Returns (trace, phases dict)."""
    r = W3dCpuRun()
    r.N[:] = list(p.N)
    r.L[:] = list(p.L)
    r.tau, r.K = p.tau, p.K
    r.periodic[:] = [int(x) for x in p.periodic]
    r.threads = threads
    r.dims[:] = list(dims)
    r.every = every
    r.verbose = int(verbose)
    trace = np.empty(2 * (p.K + 1), dtype=np.float64)
    ph = np.empty(5, dtype=np.float64)
    rc = lib().w3d_cpu_solve(C.byref(r), _p(trace), _p(ph))
    if rc != 0:
        raise RuntimeError(f"w3d_cpu_solve failed with code {rc} (blocks too thin for dims={tuple(dims)}?)")
    tr = trace.reshape(-1, 2)
    out: List[Tuple[int, float, float, float]] = [
        (n, n * p.tau, float(tr[n, 0]), float(tr[n, 1])) for n in range(1, p.K + 1) if tr[n, 0] >= 0]
    phases = dict(zip(("total", "init", "compute", "boundary", "exchange"), map(float, ph)))
    return out, phases
