"""Multi-GPU: fused P2P paths vs NCCL baseline vs the CPU solver (digit-level trace equality).

Runs only where >= 2 GPUs are visible (``gpurun --gpus 2/4/8``).  One process per GPU.
Reference parity: 1-GPU and 2-GPU traces identical (report.pdf p.15-16 §4.3).

All cases of one world size run inside ONE spawned process group (process start-up, NCCL
and symmetric-memory rendezvous cost seconds each), and every test then checks its own
case against the native CPU solver's trace.  The cases cover:

  * the three exchange paths: ``nccl`` (pack -> send/recv -> unpack), ``fused`` (one pass per
    step, in-kernel P2P halos), ``pair`` (two steps per pass; wide-halo protocol when one
    axis has neighbours, fix-up protocol otherwise);
  * protocol stress: a rotating rank is delayed on the device before every pass (and, in the
    fix-up protocol, another rank before the fix-up launch), so a flag that is released or
    consumed too early shows up as a wrong trace;
  * geometries where a tile ends exactly one layer short of a face (local extent 1 mod tile
    size; a z chunk of one plane) — there the level-1 ring of a tile reads a ghost layer
    although the tile owns no edge node (the stale-ghost race fixed in round 2).
"""
import functools

import pytest
import torch

from mpi_cuda_b200 import WaveProblem
from mpi_cuda_b200.ops import cpu
from mpi_cuda_b200.parallel.launch import spawn

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]

L_BOX = 2.0


def case(mode, dims, bc, N, K, graph=False, delay=False, zchunk=0, edges=True):
    """edges=False: blocks with neighbours along two or three axes use the fix-up protocol instead of the
    wide halo with edge lines."""
    return dict(mode=mode, dims=tuple(dims), bc=bc, N=tuple(N), K=K, graph=graph, delay=delay, zchunk=zchunk,
                edges=edges)


def _key(c):
    return (c["mode"], c["dims"], c["bc"], c["N"], c["K"], c["graph"], c["delay"], c["zchunk"], c["edges"])


SMALL = (72, 60, 52)
CASES = {
    2: [case(m, d, bc, SMALL, 8) for m in ("nccl", "fused", "pair")
        for d, bc in [((1, 1, 2), "ddd"), ((1, 2, 1), "ddp"), ((2, 1, 1), "ddd"), ((1, 1, 2), "ddp")]]
       + [case(m, (1, 1, 2), "ddd", (96, 64, 64), 10, graph=True) for m in ("fused", "pair")]
       + [case("pair", (1, 1, 2), "ddd", SMALL, 7)]
       # delayed ranks
       + [case("fused", (1, 1, 2), "ddp", SMALL, 12, delay=True),
          case("pair", (1, 1, 2), "ddp", SMALL, 12, delay=True),        # wide halo
          case("pair", (1, 2, 1), "ddp", SMALL, 12, delay=True),        # two axes: wide halo with edge lines
          case("pair", (1, 2, 1), "ddp", SMALL, 12, delay=True, edges=False),   # two axes: fix-up protocol
          case("pair", (2, 1, 1), "dpp", SMALL, 8, delay=True),         # three axes, x split: edges to the other GPU and to itself
          case("pair", (2, 1, 1), "ddd", SMALL, 9, delay=True)]         # wide halo across x
       # tiles that end one layer short of the face, delayed ranks
       + [case("pair", (1, 2, 1), "ddd", (40, 67, 30), 8, delay=True),      # n_y = 33 + 33
          case("pair", (1, 2, 1), "ddp", (40, 67, 30), 8, delay=True),      # same with a second axis (edge lines)
          case("pair", (1, 2, 1), "ddp", (40, 67, 30), 8, delay=True, edges=False),   # same, fix-up protocol
          case("pair", (1, 2, 1), "ddd", (64, 514, 24), 6, delay=True),     # 513 -> 257 + 256
          case("pair", (2, 1, 1), "ddd", (131, 30, 30), 8, delay=True),     # n_x = 65 + 65
          case("pair", (2, 1, 1), "dpd", (131, 30, 30), 8, delay=True),     # same with a second axis (edge lines)
          case("pair", (2, 1, 1), "dpd", (131, 30, 30), 8, delay=True, edges=False),  # same, fix-up protocol
          case("pair", (1, 1, 2), "ddd", (40, 40, 131), 8, delay=True, zchunk=64),   # last chunk = 1 plane
          case("pair", (1, 1, 2), "pdd", (40, 40, 131), 8, delay=True, zchunk=64),   # same with a second axis (edge lines)
          case("pair", (1, 1, 2), "pdd", (40, 40, 131), 8, delay=True, zchunk=64, edges=False)],  # same, fix-up protocol
    4: [case("pair", d, "dpd", (80, 72, 64), 8) for d in [(1, 2, 2), (2, 2, 1), (1, 1, 4), (1, 4, 1)]]
       + [case("fused", d, "dpd", (80, 72, 64), 8) for d in [(1, 2, 2), (2, 2, 1)]]
       + [case("pair", (1, 1, 4), "ddd", (80, 72, 64), 9, delay=True),
          case("pair", (1, 1, 4), "ddp", (80, 72, 64), 8, delay=True),
          case("pair", (1, 2, 2), "ddd", (80, 72, 64), 8, delay=True),
          case("pair", (1, 2, 2), "ddd", (80, 72, 64), 8, delay=True, edges=False),
          case("pair", (2, 2, 1), "ddp", (80, 72, 64), 9, delay=True),
          case("pair", (1, 2, 2), "ppp", (80, 72, 64), 8, graph=True)],
    8: [case("pair", (1, 1, 8), "ddd", (96, 80, 128), 10, delay=True),
        case("pair", (1, 1, 8), "ddp", (96, 80, 128), 10, graph=True),
        case("pair", (2, 2, 2), "ddd", (96, 80, 64), 8, delay=True),
        case("pair", (1, 2, 4), "ddd", (96, 80, 64), 8),
        case("pair", (1, 2, 4), "pdp", (96, 80, 64), 8, delay=True),
        case("pair", (2, 2, 2), "ddd", (96, 80, 64), 8, delay=True, edges=False),
        case("pair", (2, 2, 2), "ppp", (96, 80, 64), 9, graph=True),
        case("fused", (1, 1, 8), "ddd", (96, 80, 128), 10),
        case("fused", (2, 2, 2), "ppp", (96, 80, 64), 8),
        case("nccl", (1, 2, 4), "ddd", (96, 80, 64), 8)],
}


def _run_case(rank, world, c):
    import torch
    from mpi_cuda_b200.ops import cuda
    from mpi_cuda_b200.parallel.exchange import CudaBlockOps, ExchangeSolver
    from mpi_cuda_b200.parallel.fused import FusedSolver
    dev = torch.device("cuda", rank)
    p = WaveProblem(c["N"], L_BOX, 1e-3, c["K"], bc=c["bc"])
    if c["mode"] == "nccl":
        s = ExchangeSolver(p, c["dims"], lambda pp, b: CudaBlockOps(pp, b, dev), every=2, backend_name="nccl")
        return s.run().trace, "nccl"
    pair = c["mode"] == "pair"
    s = FusedSolver(p, c["dims"], dev, every=2, use_graph=c["graph"], spin_timeout_ms=20000, pair=pair,
                    zchunk=c["zchunk"], wide_edges=c["edges"])
    if not c["delay"]:
        r = s.run()
        r2 = s.run()
        assert r.trace == r2.trace
        return r.trace, s.pair_mode
    # protocol stress: a rotating rank is ~1.5 ms late (device-side sleep) before every launch
    K = c["K"]
    s.init()
    n, it = 0, 0
    while n < K:
        if rank == it % world:
            torch.cuda._sleep(3_000_000)
        if pair and n + 2 <= K:
            late = (it + 1) % world
            n = s.pair_advance(n, before_fixup=(lambda: torch.cuda._sleep(3_000_000)) if rank == late else None)
        else:
            n += 1
            s.step(n, n % 2 == 0)
        it += 1
    s.barrier()
    s.check_status()
    return s.gather_trace(), s.pair_mode


def _batch(rank, world, cases):
    out = {}
    for c in cases:
        try:
            out[_key(c)] = _run_case(rank, world, c)
        except Exception as e:          # keep going: one broken case must not hide the others
            import traceback
            out[_key(c)] = ("__error__", traceback.format_exc() + repr(e))
    return out


@functools.lru_cache(maxsize=None)
def _results(world):
    return spawn(_batch, world, (CASES[world],), backend="nccl")


@functools.lru_cache(maxsize=None)
def _base(N, bc, K):
    return cpu.solve(WaveProblem(N, L_BOX, 1e-3, K, bc=bc), every=2, threads=4)[0]


def _check(world, c):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    base = _base(c["N"], c["bc"], c["K"])
    for rank_out in _results(world):
        got = rank_out[_key(c)]
        assert got[0] != "__error__", got[1]
        trace, mode = got
        assert [r[0] for r in trace] == [r[0] for r in base]
        for a, b in zip(trace, base):
            assert a[2] == b[2], (a, b)                       # max error: digit-identical
            assert abs(a[3] - b[3]) <= 1e-12 * b[3]           # L2: summation order differs per decomposition
        if c["mode"] == "pair":
            from mpi_cuda_b200.ops import cuda
            p = WaveProblem(c["N"], L_BOX, 1e-3, c["K"], bc=c["bc"])
            axes = sum(1 for d in range(3) if c["dims"][d] > 1 or p.periodic[d])
            assert mode == (cuda.PAIR_WIDE if (axes == 1 or c["edges"]) else cuda.PAIR_FIXUP)


def _id(c):
    return "-".join([c["mode"], "x".join(map(str, c["dims"])), c["bc"], "x".join(map(str, c["N"])), f"K{c['K']}"]
                    + (["graph"] if c["graph"] else []) + (["delay"] if c["delay"] else [])
                    + ([f"zc{c['zchunk']}"] if c["zchunk"] else []) + ([] if c["edges"] else ["fixup"]))


@pytest.mark.parametrize("c", CASES[2], ids=_id)
def test_two_gpus(c):
    _check(2, c)


@pytest.mark.parametrize("c", CASES[4], ids=_id)
def test_four_gpus(c):
    _check(4, c)


@pytest.mark.parametrize("c", CASES[8], ids=_id)
def test_eight_gpus(c):
    _check(8, c)


def test_native_mpigpu_two_gpus():
    """One process, two GPUs, peer access, in-kernel halos: `mpigpu -np 2` == 1 GPU trace."""
    import os
    import subprocess
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "bin", "mpigpu")
    one = subprocess.run([exe, "160", "0.001", "20"], capture_output=True, text=True, check=True).stdout
    two = subprocess.run([exe, "-np", "2", "160", "0.001", "20"], capture_output=True, text=True, check=True).stdout
    a = [l for l in one.splitlines() if l.startswith("Step ")]
    b = [l for l in two.splitlines() if l.startswith("Step ")]
    assert len(a) == 10 and a == b          # digit-identical lines, like report.pdf p.15-16
    assert "and 2 processes" in two.splitlines()[0]
    # block partition (report.pdf p.4 §1): split along y instead of z, and along x
    for dims in (["1", "2", "1"], ["2", "1", "1"]):
        blk = subprocess.run([exe, "--dims", *dims, "160", "0.001", "20"], capture_output=True, text=True, check=True).stdout
        assert [l for l in blk.splitlines() if l.startswith("Step ")] == a
    assert "exchange 0.000000" not in two       # the exchange column is measured, not a constant


def test_native_mpigpu_blocks_four_and_eight_gpus():
    """`mpigpu --dims P Q R`: 3-D blocks in the Python-free executable, digit-identical to one GPU."""
    import os
    import subprocess
    n = torch.cuda.device_count()
    if n < 4:
        pytest.skip("needs 4 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "bin", "mpigpu")
    one = subprocess.run([exe, "128", "0.001", "10", "--bc", "dpd", "--L", "2"], capture_output=True, text=True, check=True).stdout
    a = [l for l in one.splitlines() if l.startswith("Step ")]
    cases = [["1", "2", "2"], ["2", "2", "1"]] + ([["2", "2", "2"], ["1", "2", "4"], ["1", "1", "8"]] if n >= 8 else [])
    for dims in cases:
        out = subprocess.run([exe, "--dims", *dims, "128", "0.001", "10", "--bc", "dpd", "--L", "2"],
                             capture_output=True, text=True, check=True).stdout
        assert [l for l in out.splitlines() if l.startswith("Step ")] == a, dims
