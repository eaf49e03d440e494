"""sm_100a kernels vs the native CPU solver (bit-exact) and the plain-PyTorch
restatement (tolerance).  Needs a B200: run under gpurun with ``-m gpu``.

Reference parity: "error of MPI+GPU fully matches the sequential version"
(report.pdf p.16 §4.3) and the published 512^3 trace (p.15-16).
"""
import math

import numpy as np
import pytest
import torch

from mpi_cuda_b200 import WaveProblem
from mpi_cuda_b200.models import analytic, wave_torch
from mpi_cuda_b200.ops import cpu
from mpi_cuda_b200.parallel.decomp import local_tables, plan_rank
from mpi_cuda_b200.utils import report

pytestmark = pytest.mark.gpu


def _dev():
    return torch.device("cuda", 0)


def _cuda_ops(p, variant=-1, zchunk=0, kernel="tma"):
    from mpi_cuda_b200.parallel.exchange import CudaBlockOps
    blk = plan_rank(p, (1, 1, 1), 0)
    return CudaBlockOps(p, blk, _dev(), variant=variant, zchunk=zchunk, kernel=kernel), blk


def _cpu_fields(p, steps):
    """Native CPU run on the same (unpadded) block; returns list of fields per level."""
    blk = plan_rank(p, (1, 1, 1), 0)
    b = cpu.make_block(blk.n)
    tab = local_tables(p, blk)
    A = np.zeros(cpu.block_shape(b)); B = np.zeros_like(A)
    cpu.init(A, b, *tab)
    f = [A, B]
    c = p.coeffs()
    out = [A[1:-1, 1:-1, 1:-1].copy()]
    for n in range(1, steps + 1):
        s = n - 1
        for d in range(3):
            if p.periodic[d]:
                cpu.wrap(f[s & 1], b, d)
        a, bb, cs = (1.0, 0.0, 0.5) if n == 1 else (2.0, -1.0, 1.0)
        cpu.step(f[(s & 1) ^ 1], f[s & 1], b, a, bb, [cs * v for v in c])
        out.append(f[(s & 1) ^ 1][1:-1, 1:-1, 1:-1].copy())
    return out, tab, b


def test_library_loads_and_reports_version():
    from mpi_cuda_b200.ops import cuda
    assert b"sm_100a" in cuda.lib().w3d_version()
    assert torch.cuda.get_device_capability(0)[0] == 10


def test_init_kernel_matches_tables():
    p = WaveProblem((37, 21, 18), 1.0, 1e-3, 2)
    ops, blk = _cuda_ops(p)
    ops.init()
    torch.cuda.synchronize()
    sx, sy, sz = local_tables(p, blk)
    want = sz[:, None, None] * sy[None, :, None] * sx[None, None, :]
    from mpi_cuda_b200.ops.cuda import XOFF
    got = ops.f[0].cpu().numpy()
    assert np.array_equal(got[:, :, XOFF: XOFF + blk.n[0] + 2], sx[None, None, :] * (sy[None, :, None] * sz[:, None, None]))
    assert np.allclose(got[:, :, XOFF: XOFF + blk.n[0] + 2], want, rtol=0, atol=1e-15)
    assert np.count_nonzero(got[:, :, :XOFF]) == 0
    assert float(ops.f[1].abs().max()) == 0.0


@pytest.mark.parametrize("kernel,variant", [("box", -1), ("tma", 0), ("tma", 1), ("tma", 2), ("tma", 3), ("tma", 4), ("tma", 5)])
@pytest.mark.parametrize("N", [(33, 19, 11), (130, 37, 45), (257, 40, 23)])
def test_step_kernels_bit_exact_vs_cpu(kernel, variant, N):
    p = WaveProblem(N, 1.0, 1e-3, 5)
    ops, blk = _cuda_ops(p, variant=variant, kernel=kernel)
    ref, tab, b = _cpu_fields(p, p.K)
    ops.init()
    ct = analytic.time_table(p)
    for n in range(1, p.K + 1):
        a, bb, cs = (1.0, 0.0, 0.5) if n == 1 else (2.0, -1.0, 1.0)
        ops.step(n - 1, a, bb, cs, float(ct[n]), True, n)
        got = ops.field(n & 1).cpu().numpy()
        assert np.array_equal(got, ref[n]), f"level {n}: max diff {np.max(np.abs(got - ref[n]))}"
        mx, s2 = (float(v) for v in ops.local_error(n).cpu())
        # epilogue vs CPU error of the same field
        full = np.zeros(cpu.block_shape(b)); full[1:-1, 1:-1, 1:-1] = ref[n]
        cmx, cs2 = cpu.error(full, b, *tab, float(ct[n]))
        assert mx == cmx
        assert abs(s2 - cs2) <= 1e-12 * cs2
    assert int(ops.status[0]) == 0


@pytest.mark.parametrize("zchunk", [1, 3, 7, 64])
def test_tma_zchunk_seams(zchunk):
    p = WaveProblem((140, 50, 30), 1.0, 1e-3, 3)
    ops, _ = _cuda_ops(p, variant=0, zchunk=zchunk)
    ref, _, _ = _cpu_fields(p, p.K)
    ops.init()
    ct = analytic.time_table(p)
    for n in range(1, p.K + 1):
        a, bb, cs = (1.0, 0.0, 0.5) if n == 1 else (2.0, -1.0, 1.0)
        ops.step(n - 1, a, bb, cs, float(ct[n]), False, n)
    assert np.array_equal(ops.field(p.K & 1).cpu().numpy(), ref[p.K])


def test_step_vs_plain_pytorch_reference():
    """Numerics against the whole-array PyTorch restatement of the same op."""
    p = WaveProblem(48, 1.0, 1e-3, 6)
    from mpi_cuda_b200.parallel.fused import FusedSolver
    s = FusedSolver(p, (1, 1, 1), _dev(), every=1, use_graph=False)
    res = s.run()
    ref_trace, ref_field = wave_torch.solve(p, every=1, return_field=True)
    got = s.final_field().cpu()
    assert torch.max(torch.abs(got - ref_field[1:-1, 1:-1, 1:-1])) < 5e-15
    for a, b in zip(res.trace, ref_trace):
        assert abs(a[2] - b[2]) <= 5e-15 and abs(a[3] - b[3]) <= 1e-8 * b[3]


def test_standalone_error_kernel_matches_epilogue():
    from mpi_cuda_b200.ops import cuda
    p = WaveProblem((70, 33, 29), 1.0, 1e-3, 3)
    ops, blk = _cuda_ops(p, variant=1)
    ops.init()
    ct = analytic.time_table(p)
    for n in range(1, 4):
        a, bb, cs = (1.0, 0.0, 0.5) if n == 1 else (2.0, -1.0, 1.0)
        ops.step(n - 1, a, bb, cs, float(ct[n]), True, n)
    out = torch.zeros(2, dtype=torch.float64, device=_dev())
    ws = torch.zeros(int(cuda.lib().w3d_error_workspace_bytes()), dtype=torch.uint8, device=_dev())
    for _ in range(2):   # twice: the kernel must leave its ticket ready for reuse
        cuda.error_norms(ops.f[3 & 1], blk.n, ops.tab, float(ct[3]), out, ws)
        mx, s2 = (float(v) for v in out.cpu())
        emx, es2 = (float(v) for v in ops.local_error(3).cpu())
        assert mx == emx and abs(s2 - es2) <= 1e-12 * es2


def test_pack_unpack_matches_cpu():
    p = WaveProblem((21, 14, 9), 1.0, 1e-3, 1)
    ops, blk = _cuda_ops(p)
    g = torch.Generator().manual_seed(1)
    host = torch.randn(ops.f[0].shape, dtype=torch.float64, generator=g)
    ops.f[0].copy_(host)
    from mpi_cuda_b200.ops.cuda import XOFF, pitch_for
    b = cpu.make_block(blk.n, xoff=XOFF, pitch=pitch_for(blk.n[0]))
    for axis in range(3):
        for side in range(2):
            buf = torch.empty(blk.face_elems(axis), dtype=torch.float64, device=_dev())
            ops.plan.pack_face(0, axis, side, buf, ghost=False)
            want = cpu.pack(host.numpy(), b, axis, side, ghost=False)
            assert np.array_equal(buf.cpu().numpy(), want)
            ops.f[1].zero_()
            ops.plan.unpack_face(1, axis, side, buf, ghost=True)
            chk = np.zeros(host.shape)
            cpu.unpack(chk, b, axis, side, want, ghost=True)
            assert np.array_equal(ops.f[1].cpu().numpy(), chk)     # exactly the ghost layer, nothing else


def test_wrap_axis_matches_cpu():
    p = WaveProblem((13, 11, 7), 1.0, 1e-3, 1)
    ops, blk = _cuda_ops(p)
    from mpi_cuda_b200.ops.cuda import XOFF, pitch_for
    b = cpu.make_block(blk.n, xoff=XOFF, pitch=pitch_for(blk.n[0]))
    g = torch.Generator().manual_seed(5)
    for axis in range(3):
        host = torch.randn(ops.f[0].shape, dtype=torch.float64, generator=g)
        ops.f[0].copy_(host)
        ops.plan.wrap_axis(0, axis)
        want = host.numpy().copy()
        cpu.wrap(want, b, axis)
        got = ops.f[0].cpu().numpy()
        xs = slice(XOFF, XOFF + blk.n[0] + 2)
        assert np.array_equal(got[:, :, xs], want[:, :, xs])


@pytest.mark.parametrize("bc", ["ddp", "ppp", "pdd"])
def test_fused_periodic_self_wrap_matches_cpu(bc):
    """Periodic axes on one GPU go through the same in-kernel peer-store + flag path."""
    from mpi_cuda_b200.parallel.fused import FusedSolver
    p = WaveProblem((40, 36, 28), 2.0, 1e-3, 8, bc=bc)
    s = FusedSolver(p, (1, 1, 1), _dev(), every=2, use_graph=False, pair=False)
    res = s.run()
    base, _ = cpu.solve(p, every=2)
    for a, b in zip(res.trace, base):
        assert a[2] == b[2]
        assert abs(a[3] - b[3]) <= 1e-12 * b[3]
    ref, _, _ = _cpu_fields(p, p.K)
    assert np.array_equal(s.field(p.K & 1).cpu().numpy(), ref[p.K])


@pytest.mark.parametrize("pair", [False, True])
def test_fused_graph_replay_equals_eager(pair):
    from mpi_cuda_b200.parallel.fused import FusedSolver
    p = WaveProblem(64, 1.0, 1e-3, 10)
    a = FusedSolver(p, (1, 1, 1), _dev(), every=2, use_graph=False, pair=pair).run()
    g = FusedSolver(p, (1, 1, 1), _dev(), every=2, use_graph=True, pair=pair)
    r1 = g.run(); r2 = g.run()
    assert a.trace == r1.trace == r2.trace
    ora = analytic.oracle_trace(p, 2)
    for x, o in zip(r1.trace, ora):
        assert abs(x[2] - o[2]) <= 1e-13 + 1e-6 * o[2]


def test_streaming_session_matches_batch():
    from mpi_cuda_b200.parallel.fused import FusedSolver, StreamingSession
    p = WaveProblem(56, 1.0, 1e-3, 6)
    s = FusedSolver(p, (1, 1, 1), _dev(), every=1, use_graph=False, pair=False)
    batch = s.run().trace
    sess = StreamingSession(s)
    sess.start()
    got = [sess.step(n) for n in range(1, p.K + 1)]
    sess.close()
    for (mx, s2), b in zip(got, batch):
        assert mx == b[2] and abs(math.sqrt(s2 / p.n_unknowns) - b[3]) <= 1e-15 * b[3]


@pytest.mark.parametrize("pair", [False, True])
def test_watchdog_trips_instead_of_hanging(pair):
    """A halo flag that never arrives must end in an error code, not a hung GPU."""
    from mpi_cuda_b200.parallel.fused import FusedSolver
    # (pair: two periodic axes = the fix-up protocol, whose pair pass waits for a flag that only a
    # previous fix-up launch releases; a wide-halo plan on ONE GPU publishes to itself at kernel entry)
    p = WaveProblem(48, 2.0, 1e-3, 4, bc="dpp" if pair else "ddp")
    s = FusedSolver(p, (1, 1, 1), _dev(), every=1, use_graph=False, spin_timeout_ms=30, pair=pair, wide_edges=False)
    s.init()
    if pair:
        s.pair_advance(2)         # level 2 was never produced: its flag cannot arrive
    else:
        s.step(2)                 # level 1 was never produced: its flag cannot arrive
    torch.cuda.synchronize()
    with pytest.raises(RuntimeError, match="watchdog"):
        s.check_status()
    s.run()                       # the solver is usable again after a re-init
    s.check_status()


@pytest.mark.parametrize("pair,at,bc,edges", [(False, 6, "dpd", True), (True, 6, "dpd", True), (True, 5, "dpd", True),
                                              (True, 4, "ppd", False), (True, 5, "ppd", True)])
def test_checkpoint_restart_is_bit_exact(tmp_path, pair, at, bc, edges):
    """Dump after level `at`, restore into a fresh solver, finish: same field and trace as one run.

    Pair mode: the live levels sit in any two of the four arrays (the dump records which), an odd
    `at` leaves the newest level produced by a single step, and "ppd" runs the fix-up protocol
    (edges=False) or the wide halo with edge lines (edges=True)."""
    from mpi_cuda_b200.parallel.fused import FusedSolver
    p = WaveProblem((60, 44, 36), 2.0, 1e-3, 10, bc=bc)
    a = FusedSolver(p, (1, 1, 1), _dev(), every=1, use_graph=False, pair=pair, wide_edges=edges)
    full = a.run().trace
    want = a.final_field().clone()
    b = FusedSolver(p, (1, 1, 1), _dev(), every=1, use_graph=False, pair=pair, wide_edges=edges)
    b.init()
    b.run_range(1, at)
    ck = b.save_checkpoint(str(tmp_path / "ck"), at)
    assert ck.endswith(".rank0.pt")
    c = FusedSolver(p, (1, 1, 1), _dev(), every=1, use_graph=False, pair=pair, wide_edges=edges)
    level = c.load_checkpoint(str(tmp_path / "ck"))
    assert level == at
    c.run_range(level + 1, p.K)
    c.barrier()
    c.check_status()
    assert torch.equal(c.final_field(), want)
    got = c.gather_trace()
    # (an odd `at` moves the lone single step from the end of the run to the middle: the max is
    # digit-identical, the sum of squares is folded in a different order)
    for x, y in zip(got[at:], full[at:]):
        assert x[0] == y[0] and x[2] == y[2] and abs(x[3] - y[3]) <= 1e-12 * y[3]


def test_profile_steps_and_chrome_trace(tmp_path):
    import json
    from mpi_cuda_b200.parallel.fused import FusedSolver
    from mpi_cuda_b200.utils.timing import write_chrome_trace
    p = WaveProblem(96, 1.0, 1e-3, 5)
    s = FusedSolver(p, (1, 1, 1), _dev(), every=1, use_graph=False)
    ms = s.profile_steps()
    assert len(ms) == 5 and all(0 < v < 50 for v in ms)
    out = tmp_path / "trace.json"
    write_chrome_trace(str(out), ms)
    assert len(json.load(open(out))["traceEvents"]) == 5


@pytest.mark.parametrize("K", [6, 7])
def test_streaming_session_in_pair_mode(K):
    from mpi_cuda_b200.parallel.fused import FusedSolver, StreamingSession
    p = WaveProblem(56, 2.0, 1e-3, K, bc="ddp")
    base, _ = cpu.solve(p, every=1)
    s = FusedSolver(p, (1, 1, 1), _dev(), every=1, use_graph=False, pair=True)
    sess = StreamingSession(s)
    sess.start()
    for n in range(1, K + 1):
        sess.step_async(n)
    got = [sess.result(n) for n in range(1, K + 1)]
    sess.close()
    for (mx, s2), b in zip(got, base):
        assert mx == b[2] and abs(math.sqrt(s2 / p.n_unknowns) - b[3]) <= 1e-12 * b[3]


def test_golden_trace_512_on_gpu():
    """Published trace of the reference (report.pdf p.15-16 §4.3), on one B200."""
    from mpi_cuda_b200.parallel.fused import FusedSolver
    p = WaveProblem(512, 1.0, 1e-3, 20)
    s = FusedSolver(p, (1, 1, 1), _dev(), every=2, use_graph=True, pair=False)
    res = s.run()
    assert len(res.trace) == 10
    for (n, t, mx, l2), (gn, gt, gmx, gl2) in zip(res.trace, report.GOLDEN_512):
        assert n == gn and abs(mx - gmx) <= 5e-15 and abs(l2 - gl2) <= 1e-6 * gl2
    assert res.lines()[-1][:60] == "Step 20, t = 0.020000, Max Error = 3.960129e-09, L2 Error = 1.40422"[:60]


def test_native_mpigpu_executable_matches_python_path():
    """bin/mpigpu: the Python-free equivalent of `mpirun -np 1 ./mpigpu-1 N tau K 1` (report.pdf p.15)."""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "bin", "mpigpu")
    if not os.path.exists(exe):
        from mpi_cuda_b200.build import build_tools
        build_tools()
    out = subprocess.run([exe, "128", "0.001", "20", "1"], capture_output=True, text=True, check=True).stdout
    lines = [l for l in out.splitlines() if l.startswith("Step ")]
    assert out.splitlines()[0] == "Running test with grid size 128^3" and len(lines) == 10
    base, _ = cpu.solve(WaveProblem(128, 1.0, 1e-3, 20), every=2)
    for l, b in zip(lines, base):
        n, t, mx, l2 = report.parse_step(l)
        assert n == b[0] and abs(mx - b[2]) <= 1e-6 * b[2] and abs(l2 - b[3]) <= 1e-6 * b[3]
    per = subprocess.run([exe, "96", "0.001", "10", "--bc", "ddp", "--L", "2"], capture_output=True, text=True, check=True).stdout
    base, _ = cpu.solve(WaveProblem(96, 2.0, 1e-3, 10, bc="ddp"), every=2)
    got = [report.parse_step(l) for l in per.splitlines() if l.startswith("Step ")]
    assert len(got) == 5 and abs(got[-1][2] - base[-1][2]) <= 1e-6 * base[-1][2]
    assert "wide halo" in per
    # two periodic axes: fix-up protocol; odd K: the last step goes alone
    per2 = subprocess.run([exe, "96", "0.001", "9", "--bc", "pdp", "--L", "2", "--every", "1"], capture_output=True, text=True, check=True).stdout
    base2, _ = cpu.solve(WaveProblem(96, 2.0, 1e-3, 9, bc="pdp"), every=1)
    got2 = [report.parse_step(l) for l in per2.splitlines() if l.startswith("Step ")]
    assert "wide halo" in per2 and len(got2) == 9      # edge lines to the diagonal neighbour (the GPU itself)
    for g, b in zip(got2, base2):
        assert abs(g[2] - b[2]) <= 1e-6 * b[2] and abs(g[3] - b[3]) <= 1e-6 * b[3]
    per3 = subprocess.run([exe, "96", "0.001", "9", "--bc", "pdp", "--L", "2", "--every", "1", "--fixup"], capture_output=True, text=True, check=True).stdout
    assert "fix-up" in per3 and [l for l in per3.splitlines() if l.startswith("Step ")] == [l for l in per2.splitlines() if l.startswith("Step ")]


@pytest.mark.parametrize("pair_variant", [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10])
@pytest.mark.parametrize("N,K", [((70, 45, 37), 6), ((130, 66, 50), 7), ((257, 40, 23), 4)])
def test_pair_kernel_bit_exact_vs_cpu(N, K, pair_variant):
    """Temporal blocking (two steps per pass) must give exactly the fields of K single steps."""
    from mpi_cuda_b200.parallel.fused import FusedSolver
    p = WaveProblem(N, 1.0, 1e-3, K)
    s = FusedSolver(p, (1, 1, 1), _dev(), every=1, use_graph=False, pair=True, pair_variant=pair_variant)
    res = s.run()
    ref, tab, b = _cpu_fields(p, K)
    got = s.final_field().cpu().numpy()
    assert np.array_equal(got, ref[K]), f"max diff {np.max(np.abs(got - ref[K]))}"
    base, _ = cpu.solve(p, every=1)
    assert len(res.trace) == K
    for a, c in zip(res.trace, base):
        assert a[2] == c[2] and abs(a[3] - c[3]) <= 1e-12 * c[3]


@pytest.mark.parametrize("bc,edges", [("ddp", True), ("dpd", True), ("pdd", True), ("ppp", True), ("dpp", True), ("pdp", True),
                                      ("ppd", True), ("ppp", False), ("dpp", False)])
@pytest.mark.parametrize("K", [6, 7])
def test_pair_mode_with_periodic_faces_matches_cpu(bc, edges, K):
    """Two-step passes with in-kernel halo exchange (periodic axes wrap onto the GPU itself).

    One periodic axis: wide halo (one launch, one handshake per two steps).  Two or three: wide halo with the
    edge lines of the diagonal neighbours (here the GPU itself), or — edges=False — the fix-up protocol."""
    from mpi_cuda_b200.ops import cuda
    from mpi_cuda_b200.parallel.fused import FusedSolver
    p = WaveProblem((70, 45, 37), 2.0, 1e-3, K, bc=bc)
    s = FusedSolver(p, (1, 1, 1), _dev(), every=1, use_graph=False, pair=True, wide_edges=edges)
    assert s.pair_mode == (cuda.PAIR_WIDE if (bc.count("p") == 1 or edges) else cuda.PAIR_FIXUP)
    res = s.run()
    s.check_status()
    ref, _, _ = _cpu_fields(p, K)
    got = s.final_field().cpu().numpy()
    assert np.array_equal(got, ref[K]), f"max diff {np.max(np.abs(got - ref[K]))}"
    base, _ = cpu.solve(p, every=1)
    for a, c in zip(res.trace, base):
        assert a[2] == c[2] and abs(a[3] - c[3]) <= 1e-12 * c[3]
    r2 = s.run()
    assert r2.trace == res.trace


def test_pair_kernel_golden_trace_and_graph():
    from mpi_cuda_b200.parallel.fused import FusedSolver
    p = WaveProblem(512, 1.0, 1e-3, 20)
    s = FusedSolver(p, (1, 1, 1), _dev(), every=2, use_graph=True, pair=True)
    res = s.run()
    for (n, t, mx, l2), (gn, gt, gmx, gl2) in zip(res.trace, report.GOLDEN_512):
        assert n == gn and abs(mx - gmx) <= 5e-15 and abs(l2 - gl2) <= 1e-6 * gl2


@pytest.mark.parametrize("N,K,bc", [((9, 7, 5), 3, "ddd"), ((65, 33, 6), 4, "ddp"), ((12, 70, 9), 1, "ppp"),
                                    ((4, 4, 4), 2, "ddd"), ((200, 9, 140), 5, "dpd")])
@pytest.mark.parametrize("pair", [False, True])
def test_small_and_odd_blocks(N, K, bc, pair):
    """Blocks smaller than a tile, one-plane-thin chunks, K = 1, every axis periodic."""
    from mpi_cuda_b200.parallel.fused import FusedSolver
    p = WaveProblem(N, 3.0, 1e-3, K, bc=bc)
    s = FusedSolver(p, (1, 1, 1), _dev(), every=1, use_graph=False, pair=pair)
    res = s.run()
    s.check_status()
    ref, _, _ = _cpu_fields(p, K)
    assert np.array_equal(s.final_field().cpu().numpy(), ref[K])
    base, _ = cpu.solve(p, every=1)
    for a, c in zip(res.trace, base):
        assert a[2] == c[2] and abs(a[3] - c[3]) <= 1e-12 * max(c[3], 1e-300)


@pytest.mark.parametrize("N,bc,zchunk,pv", [((65, 40, 30), "pdd", 0, -1), ((40, 33, 30), "dpd", 0, -1),
                                            ((70, 66, 65), "ddp", 64, -1), ((129, 30, 33), "pdd", 16, 9),
                                            ((48, 61, 40), "dpd", 7, 8), ((30, 30, 3), "ddp", 0, -1),
                                            ((3, 40, 20), "pdd", 0, -1), ((40, 3, 20), "dpd", 1, -1)])
@pytest.mark.parametrize("K", [4, 5])
def test_wide_halo_edge_geometries(N, bc, zchunk, pv, K):
    """Wide-halo pair passes where the shipped layers straddle tiles / chunks.

    A periodic axis with n = 1 (mod tile) puts the ghost column inside a partial tile and the
    layers n-1, n into different tiles; zchunk leaves thin last chunks; 2- and 3-node-thick
    blocks make the two shipped layers of both faces overlap.  Bit-exact against the CPU."""
    from mpi_cuda_b200.ops import cuda
    from mpi_cuda_b200.parallel.fused import FusedSolver
    p = WaveProblem(N, 3.0, 1e-3, K, bc=bc)
    s = FusedSolver(p, (1, 1, 1), _dev(), every=1, use_graph=False, pair=True, zchunk=zchunk, pair_variant=pv)
    assert s.pair_mode == cuda.PAIR_WIDE
    res = s.run()
    s.check_status()
    ref, _, _ = _cpu_fields(p, K)
    got = s.final_field().cpu().numpy()
    assert np.array_equal(got, ref[K]), f"max diff {np.max(np.abs(got - ref[K]))}"
    base, _ = cpu.solve(p, every=1)
    for a, c in zip(res.trace, base):
        assert a[2] == c[2] and abs(a[3] - c[3]) <= 1e-12 * max(c[3], 1e-300)
    assert s.run().trace == res.trace          # second run on the same arrays (stale ghosts must not matter)


def test_exchange_statistics_are_measured():
    """The phase table's exchange columns come from %globaltimer stamps inside the face CTAs."""
    from mpi_cuda_b200.parallel.fused import FusedSolver
    p = WaveProblem((128, 128, 64), 2.0, 1e-3, 8, bc="ddp")
    s = FusedSolver(p, (1, 1, 1), _dev(), every=2, use_graph=True, pair=True)
    r = s.run()
    xs = r.extra["exchange_stats"]
    assert xs["launches"] == 4                      # one handshake per two steps
    assert 0 <= r.phases["halo_copy"] < r.phases["loop"] and 0 < r.phases["flag_wait"] < r.phases["loop"]
    assert r.phases["exchange"] == r.phases["flag_wait"] + r.phases["halo_copy"]
    f = FusedSolver(p, (1, 1, 1), _dev(), every=2, use_graph=False, pair=False)
    assert f.run().extra["exchange_stats"]["launches"] == 8


@pytest.mark.parametrize("N,L,want", [(512, math.pi, 6.141989e-10), (128, math.pi, 1.071891e-08)])
def test_published_delta_on_L_pi(N, L, want):
    """delta for Omega = [0,pi]^3 as published (report.pdf p.8, p.10, p.11: 1.07e-8 / 6.14e-10)."""
    from mpi_cuda_b200.parallel.fused import FusedSolver
    p = WaveProblem(N, L, 1e-3, 20)
    res = FusedSolver(p, (1, 1, 1), _dev(), every=2, use_graph=True).run()
    n, t, mx, l2 = res.trace[-1]
    assert n == 20 and abs(l2 - want) <= 2e-6 * want
    omx, ol2 = analytic.oracle_errors(p, 20)
    assert abs(mx - omx) <= 5e-15 and abs(l2 - ol2) <= 1e-6 * ol2


@pytest.mark.parametrize("N,bc,K", [((130, 65, 40), "ddd", 6), ((70, 100, 30), "ddp", 5), ((200, 129, 70), "ddd", 4),
                                    ((65, 257, 24), "pdd", 4)])
def test_cluster_variant_bit_exact(N, bc, K):
    """2-CTA thread-block clusters along y: the ring row between the two tiles comes from the partner's
    shared memory (DSMEM) after a cluster barrier instead of being recomputed.  Same bits as the CPU.

    Sizes: an exact pair of tiles; four tile rows with the last one moved back by 29 rows (the DSMEM row
    offsets depend on that overlap); several clusters; a periodic x axis (ghost column in a partial tile)."""
    from mpi_cuda_b200.parallel.fused import FusedSolver
    p = WaveProblem(N, 2.0, 1e-3, K, bc=bc)
    s = FusedSolver(p, (1, 1, 1), _dev(), every=1, use_graph=False, pair=True, pair_variant=11)
    assert "cluster" in s.describe()
    res = s.run()
    s.check_status()
    ref, _, _ = _cpu_fields(p, K)
    got = s.final_field().cpu().numpy()
    assert np.array_equal(got, ref[K]), f"max diff {np.max(np.abs(got - ref[K]))}"
    base, _ = cpu.solve(p, every=1)
    for a, c in zip(res.trace, base):
        assert a[2] == c[2] and abs(a[3] - c[3]) <= 1e-12 * c[3]


def test_cluster_variant_needs_even_tile_rows():
    from mpi_cuda_b200.ops.cuda import CudaLibError
    from mpi_cuda_b200.parallel.fused import FusedSolver
    with pytest.raises(CudaLibError, match="even number of tile rows"):
        FusedSolver(WaveProblem((70, 70, 30), 2.0, 1e-3, 2), (1, 1, 1), _dev(), pair=True, pair_variant=11)


def test_north_star_1024_L_pi_K100_matches_oracle():
    """1024^3 on [0,pi]^3, K = 100 (the CFL-stable north-star configuration; SURVEY.md App. A.4:
    Max 7.975469e-09, L2 2.823890e-09 at step 100) on ONE GPU: 4 x 8.7 GB arrays, 0.3 s of stepping."""
    from mpi_cuda_b200.parallel.fused import FusedSolver
    p = WaveProblem(1024, math.pi, 1e-3, 100)
    s = FusedSolver(p, (1, 1, 1), _dev(), every=10, use_graph=True)
    res = s.run()
    s.check_status()
    assert len(res.trace) == 10
    for n, t, mx, l2 in res.trace:
        omx, ol2 = analytic.oracle_errors(p, n)
        assert abs(mx - omx) <= 5e-15 * max(1.0, n / 20.0) and abs(l2 - ol2) <= 1e-6 * ol2
    n, t, mx, l2 = res.trace[-1]
    assert n == 100 and abs(mx - 7.975469e-09) <= 3e-14 and abs(l2 - 2.823890e-09) <= 3e-15
    del s
    torch.cuda.empty_cache()
