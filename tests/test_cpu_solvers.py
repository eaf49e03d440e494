"""Native CPU solvers vs the plain-PyTorch restatement vs the closed-form oracle.

Reference parity: cross-version agreement and thread/rank-count invariance of the
error (report.pdf p.7-11 delta columns, p.16 §4.3).
"""
import math
import os
import subprocess

import numpy as np
import pytest
import torch

from mpi_cuda_b200 import WaveProblem, solve
from mpi_cuda_b200.models import analytic, wave_torch
from mpi_cuda_b200.ops import cpu
from mpi_cuda_b200.utils import report

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _close(tr_a, tr_b, abs_max=5e-15, rel_l2=1e-9):
    assert [r[0] for r in tr_a] == [r[0] for r in tr_b]
    for a, b in zip(tr_a, tr_b):
        assert abs(a[2] - b[2]) <= abs_max + 1e-9 * abs(b[2]), (a, b)
        assert abs(a[3] - b[3]) <= rel_l2 * abs(b[3]) + 1e-18, (a, b)


@pytest.mark.parametrize("bc,L", [("ddd", 1.0), ("ddp", 2.0), ("ppp", 3.0), ("pdd", math.pi)])
def test_native_vs_torch_vs_oracle(bc, L):
    p = WaveProblem(40, L, 1e-3, 12, bc=bc)
    nat, _ = cpu.solve(p, every=2)
    ref = wave_torch.solve(p, every=2)
    _close(nat, ref)
    ora = analytic.oracle_trace(p, every=2)
    for a, o in zip(nat, ora):
        assert abs(a[2] - o[2]) <= 1e-13 + 1e-6 * o[2]
        assert abs(a[3] - o[3]) <= 1e-13 + 1e-6 * o[3]


def test_native_field_matches_torch_field():
    p = WaveProblem(24, 1.0, 1e-3, 6)
    b = cpu.make_block(p.unknowns)
    tab = tuple(np.ascontiguousarray(t) for t in analytic.axis_tables(p))
    A = np.zeros(cpu.block_shape(b)); B = np.zeros_like(A)
    cpu.init(A, b, *tab)
    c = p.coeffs()
    cpu.step(B, A, b, 1.0, 0.0, [0.5 * v for v in c])
    cur, prv = B, A
    for n in range(2, p.K + 1):
        cpu.step(prv, cur, b, 2.0, -1.0, list(c))
        cur, prv = prv, cur
    _, ref = wave_torch.solve(p, return_field=True)
    assert np.max(np.abs(cur - ref.numpy())) < 5e-15


@pytest.mark.parametrize("threads", [1, 2, 5])
def test_openmp_thread_count_invariance(threads):
    p = WaveProblem(48, 1.0, 1e-3, 10)
    base, _ = cpu.solve(p, threads=1, every=2)
    got, _ = cpu.solve(p, threads=threads, every=2)
    assert got == base     # bit-identical: fixed per-plane reduction order


@pytest.mark.parametrize("dims", [(1, 1, 2), (1, 2, 2), (2, 2, 2), (1, 1, 5), (3, 1, 2)])
@pytest.mark.parametrize("bc", ["ddd", "dpp"])
def test_rank_count_invariance_native(dims, bc):
    p = WaveProblem(36, 1.0, 1e-3, 8, bc=bc)
    base, _ = cpu.solve(p, every=2)
    got, ph = cpu.solve(p, threads=2, dims=dims, every=2)
    for a, b in zip(got, base):
        assert a[2] == b[2]                      # max error digit-identical (report.pdf p.15-16)
        assert abs(a[3] - b[3]) <= 1e-13 * b[3]
    assert set(ph) == {"total", "init", "compute", "boundary", "exchange"}


@pytest.mark.slow
def test_golden_trace_512_native_openmp():
    """The published 10-line trace (report.pdf p.15-16 §4.3), reproduced by the native solver."""
    p = WaveProblem(512, 1.0, 1e-3, 20)
    got, _ = cpu.solve(p, threads=min(8, os.cpu_count() or 1), every=2)
    assert len(got) == 10
    for (n, t, mx, l2), (gn, gt, gmx, gl2) in zip(got, report.GOLDEN_512):
        assert n == gn
        assert abs(mx - gmx) <= 5e-15
        assert abs(l2 - gl2) <= 1e-6 * gl2


def test_pack_unpack_roundtrip():
    b = cpu.make_block((5, 4, 3), xoff=3, pitch=12)
    rng = np.random.default_rng(0)
    u = rng.standard_normal(cpu.block_shape(b))
    for axis in range(3):
        for side in range(2):
            buf = cpu.pack(u, b, axis, side, ghost=False)
            assert buf.size == cpu.face_elems(b, axis)
            v = np.zeros_like(u)
            cpu.unpack(v, b, axis, side, buf, ghost=False)
            assert np.count_nonzero(v) == buf.size
            idx = 1 if side == 0 else b.n[axis]
            sl = [slice(1, b.n[2] + 1), slice(1, b.n[1] + 1), slice(b.xoff + 1, b.xoff + 1 + b.n[0])]
            sl[2 - axis] = (idx if axis else b.xoff + idx)
            assert np.array_equal(np.sort(u[tuple(sl)].ravel()), np.sort(buf))


def test_solve_api_backends_agree():
    p = WaveProblem(32, 1.0, 1e-3, 6)
    r1 = solve(p, backend="cpu")
    r2 = solve(p, backend="cpu-omp", threads=3)
    r3 = solve(p, backend="cpu-ranks", ranks=4)
    r4 = solve(p, backend="torch")
    assert r1.trace == r2.trace
    _close(r3.trace, r1.trace)
    _close(r4.trace, r1.trace)
    assert r1.updates == 31 ** 3 * 6 and r1.glups > 0


def test_native_executables_reference_cli():
    exe = os.path.join(ROOT, "bin", "wave")
    if not os.path.exists(exe):
        from mpi_cuda_b200.build import build_cpu
        build_cpu()
    out = subprocess.run([exe, "64", "0.001", "20"], capture_output=True, text=True, check=True).stdout
    lines = [l for l in out.splitlines() if l.startswith("Step ")]
    assert out.splitlines()[0] == "Running test with grid size 64^3"
    assert len(lines) == 10
    n, t, mx, l2 = report.parse_step(lines[-1])
    o = analytic.oracle_errors(WaveProblem(64, 1.0, 1e-3, 20), 20)
    assert n == 20 and abs(mx - o[0]) < 1e-6 * o[0] and abs(l2 - o[1]) < 1e-6 * o[1]
    out2 = subprocess.run([os.path.join(ROOT, "bin", "mpiomp"), "-np", "4", "64", "0.001", "20", "2"],
                          capture_output=True, text=True, check=True).stdout
    assert "and 4 processes" in out2.splitlines()[0]
    assert [l for l in out2.splitlines() if l.startswith("Step ")] == lines
