"""Error norms without any device allocation (``w3d_error_norms_noalloc``).

For callers that must not call ``cudaMalloc`` — a solver embedded in an
application that owns all device memory through its own pool, or a hook inside
another code's time loop where an allocation would synchronise the device.  The
scratch the reduction needs is a fixed array inside ``libwave3d_b200.so``
(33 024 bytes per GPU), so the block cross-section is limited to
``MAX_ROWS = ny*nz <= 262 144``; beyond that :class:`BlockTooLarge` is raised and
the regular :func:`mpi_cuda_b200.ops.cuda.error_norms` (caller-provided
workspace) has to be used.  Calls on one device share that array: issue them on
one stream, or order them yourself.  This entry point has no counterpart in the
reference.
"""
from __future__ import annotations

import ctypes as C
from typing import Sequence

import torch

from . import cuda

MAX_ROWS = 262144
SCRATCH_BYTES = 33024
E_TOOBIG = -5


class BlockTooLarge(ValueError):
    pass


def _fn():
    f = cuda.lib().w3d_error_norms_noalloc
    f.restype = C.c_int
    f.argtypes = [C.c_void_p, C.POINTER(cuda.W3dBlock), C.c_void_p, C.c_void_p, C.c_void_p, C.c_double,
                  C.c_void_p, C.c_void_p, C.c_char_p, C.c_size_t]
    return f


def error_norms_noalloc(u: torch.Tensor, n: Sequence[int], tables: Sequence[torch.Tensor], ct: float,
                        out: torch.Tensor) -> None:
    """out[0] = max |u - exact|, out[1] = sum of squares; ``out`` is a 2-element cuda float64 tensor."""
    blk = cuda.make_block(n)
    msg = C.create_string_buffer(512)
    rc = _fn()(u.data_ptr(), C.byref(blk), tables[0].data_ptr(), tables[1].data_ptr(), tables[2].data_ptr(),
               float(ct), out.data_ptr(), cuda.stream_ptr(), msg, len(msg))
    if rc == E_TOOBIG:
        raise BlockTooLarge(msg.value.decode())
    if rc != 0:
        raise cuda.CudaLibError(f"w3d_error_norms_noalloc failed ({rc}): {msg.value.decode()}")
