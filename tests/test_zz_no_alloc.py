"""``w3d_error_norms_noalloc`` — the entry point for callers that cannot allocate.

Some users of this library run inside an application that hands out all device
memory once from its own pool, or inside another solver's time loop where a
``cudaMalloc`` would stall every stream on the device.  They can give us inputs
and an output but no workspace, so the library keeps the scratch of the error
reduction in a built-in 33 024-byte array per GPU.  This file checks, through the
C ABI (ctypes, raw device pointers), that the result agrees with the regular
``w3d_error_norms`` on calls in a row — large field, small field, large field
again — and that an oversized block is refused with a readable message and leaves
the output alone.  The sequence is also the dirty-ticket case: every call leaves its
CTA count in the array's ticket word, so the next call (fewer CTAs, then more) only
finds its "last CTA" — and only then writes the output at all — if the entry point
really clears the word first; the output is pre-filled with a sentinel to see that.

Collected last on purpose (file name): a failure here cannot hide other tests
from a run that stops at the first failure.  Needs a GPU.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(n, seed):
    from mpi_cuda_b200.ops import cuda
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(seed)
    shape = cuda.field_shape(n)
    u = torch.randn(shape, dtype=torch.float64, generator=g).to(dev)
    tabs = [torch.rand(k + 2, dtype=torch.float64, generator=g).to(dev) for k in n]
    return cuda, dev, u, tabs


def _reference(cuda, dev, u, n, tabs, ct):
    """The existing entry point, with a caller-provided workspace."""
    out = torch.zeros(2, dtype=torch.float64, device=dev)
    ws = torch.zeros(int(cuda.lib().w3d_error_workspace_bytes()), dtype=torch.uint8, device=dev)
    cuda.error_norms(u, n, tabs, ct, out, ws)
    torch.cuda.synchronize()
    return [float(v) for v in out.cpu()]


def _call_c_abi(cuda, u, n, tabs, ct, out):
    """Straight through the C ABI: no Python wrapper in between."""
    fn = cuda.lib().w3d_error_norms_noalloc
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.POINTER(cuda.W3dBlock), C.c_void_p, C.c_void_p, C.c_void_p, C.c_double,
                   C.c_void_p, C.c_void_p, C.c_char_p, C.c_size_t]
    blk = cuda.make_block(n)
    msg = C.create_string_buffer(512)
    rc = fn(u.data_ptr(), C.byref(blk), tabs[0].data_ptr(), tabs[1].data_ptr(), tabs[2].data_ptr(), ct,
            out.data_ptr(), cuda.stream_ptr(), msg, len(msg))
    return rc, msg.value.decode()


def test_noalloc_matches_workspace_version_twice_in_a_row():
    first, second = (150, 67, 45), (33, 20, 9)            # 24 CTAs, then 2 CTAs, then 24 again
    cuda, dev, u1, t1 = _setup(first, 1)
    _, _, u2, t2 = _setup(second, 2)
    out = torch.full((2,), -1.0, dtype=torch.float64, device=dev)
    before = torch.cuda.memory_allocated(dev)
    for n, u, tabs, ct in ((first, u1, t1, 0.75), (second, u2, t2, -0.3), (first, u1, t1, 0.1)):
        out.fill_(-1.0)                                    # sentinel: a call that never finds its last CTA leaves it
        rc, msg = _call_c_abi(cuda, u, n, tabs, ct, out)
        assert rc == 0, msg
        torch.cuda.synchronize()
        got = [float(v) for v in out.cpu()]
        assert got[0] >= 0.0 and got[1] >= 0.0, "output not written: the ticket word was not cleared"
        want = _reference(cuda, dev, u, n, tabs, ct)
        assert got[0] == want[0]                               # max is order-independent
        assert abs(got[1] - want[1]) <= 1e-12 * want[1]        # sum: same tolerance as the other GPU tests
        # independent check of the op itself against plain PyTorch
        xo = cuda.XOFF
        core = u[1:-1, 1:-1, xo + 1: xo + 1 + n[0]]
        exact = (tabs[2][1:-1, None, None] * ct) * (tabs[0][None, None, 1:-1] * tabs[1][None, 1:-1, None])
        e = (core - exact).abs()
        assert abs(got[0] - float(e.max())) <= 1e-15 * max(1.0, got[0]) + 1e-15
        assert abs(got[1] - float((e * e).sum())) <= 1e-11 * got[1]
    del core, exact, e, want
    assert torch.cuda.memory_allocated(dev) - before < (1 << 22)   # only the test's own small tensors


def test_noalloc_rejects_blocks_beyond_the_builtin_scratch():
    from mpi_cuda_b200.ops import noalloc
    n = (8, 513, 512)                                            # ny*nz = 262 656 > 262 144
    cuda, dev, u, tabs = _setup(n, 3)
    out = torch.full((2,), -7.0, dtype=torch.float64, device=dev)
    rc, msg = _call_c_abi(cuda, u, n, tabs, 1.0, out)
    assert rc == noalloc.E_TOOBIG
    assert "262144" in msg and "w3d_error_norms()" in msg
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [-7.0, -7.0]
    with pytest.raises(noalloc.BlockTooLarge):
        noalloc.error_norms_noalloc(u, n, tabs, 1.0, out)
    # the largest allowed cross-section still works
    n_ok = (8, 512, 512)
    _, _, u2, t2 = _setup(n_ok, 4)
    noalloc.error_norms_noalloc(u2, n_ok, t2, 0.5, out)
    torch.cuda.synchronize()
    want = _reference(cuda, dev, u2, n_ok, t2, 0.5)
    got = out.cpu().tolist()
    assert got[0] == want[0] and abs(got[1] - want[1]) <= 1e-12 * want[1]
