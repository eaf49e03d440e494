"""Multi-GPU: fused P2P path vs NCCL baseline vs 1 GPU (digit-level trace equality).

Runs only where >= 2 GPUs are visible (``gpurun --gpus 2``).  One process per GPU.
Reference parity: 1-GPU and 2-GPU traces identical (report.pdf p.15-16 §4.3).
"""
import pytest
import torch

from mpi_cuda_b200 import WaveProblem
from mpi_cuda_b200.ops import cpu
from mpi_cuda_b200.parallel.launch import spawn

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _work_skew(rank, world, dims, bc, N, K):
    """Protocol stress: one rank is artificially late before every step."""
    import torch
    from mpi_cuda_b200.parallel.fused import FusedSolver
    dev = torch.device("cuda", rank)
    p = WaveProblem(N, 2.0, 1e-3, K, bc=bc)
    s = FusedSolver(p, dims, dev, every=2, use_graph=False, spin_timeout_ms=20000)
    s.init()
    for n in range(1, K + 1):
        if rank == (n % world):
            torch.cuda._sleep(3_000_000)          # ~1.5 ms of device-side delay on a rotating rank
        s.step(n, n % 2 == 0)
    s.barrier()
    s.check_status()
    return s.gather_trace()


def _work(rank, world, mode, dims, bc, N, K, graph):
    import torch
    from mpi_cuda_b200.parallel.exchange import CudaBlockOps, ExchangeSolver
    from mpi_cuda_b200.parallel.fused import FusedSolver
    dev = torch.device("cuda", rank)
    p = WaveProblem(N, 2.0, 1e-3, K, bc=bc)
    if mode in ("fused", "pair"):
        s = FusedSolver(p, dims, dev, every=2, use_graph=graph, spin_timeout_ms=5000, pair=(mode == "pair"))
        r = s.run()
        r2 = s.run()
        assert r.trace == r2.trace
        return r.trace
    s = ExchangeSolver(p, dims, lambda pp, b: CudaBlockOps(pp, b, dev), every=2, backend_name="nccl")
    return s.run().trace


def _ngpu():
    return torch.cuda.device_count()


@pytest.mark.parametrize("mode", ["nccl", "fused", "pair"])
@pytest.mark.parametrize("dims,bc", [((1, 1, 2), "ddd"), ((1, 2, 1), "ddp"), ((2, 1, 1), "ddd"), ((1, 1, 2), "ddp")])
def test_two_gpu_trace_equals_one_gpu(mode, dims, bc):
    N, K = (72, 60, 52), 8
    base, _ = cpu.solve(WaveProblem(N, 2.0, 1e-3, K, bc=bc), every=2)
    out = spawn(_work, 2, (mode, dims, bc, N, K, False), backend="nccl")
    for trace in out:
        assert [r[0] for r in trace] == [r[0] for r in base]
        for a, b in zip(trace, base):
            assert a[2] == b[2]
            assert abs(a[3] - b[3]) <= 1e-12 * b[3]


def test_two_gpu_skewed_ranks_still_agree():
    N, K = (72, 60, 52), 12
    base, _ = cpu.solve(WaveProblem(N, 2.0, 1e-3, K, bc="ddp"), every=2)
    out = spawn(_work_skew, 2, ((1, 1, 2), "ddp", N, K), backend="nccl")
    for trace in out:
        for a, b in zip(trace, base):
            assert a[2] == b[2] and abs(a[3] - b[3]) <= 1e-12 * b[3]


@pytest.mark.parametrize("mode", ["fused", "pair"])
def test_two_gpu_fused_graph(mode):
    N, K = (96, 64, 64), 10
    base, _ = cpu.solve(WaveProblem(N, 2.0, 1e-3, K), every=2)
    out = spawn(_work, 2, (mode, (1, 1, 2), "ddd", N, K, True), backend="nccl")
    for a, b in zip(out[0], base):
        assert a[2] == b[2]


@pytest.mark.parametrize("dims", [(1, 2, 2), (2, 2, 1), (1, 1, 4)])
def test_four_gpu_blocks(dims):
    if _ngpu() < 4:
        pytest.skip("needs 4 GPUs")
    N, K = (80, 72, 64), 8
    base, _ = cpu.solve(WaveProblem(N, 2.0, 1e-3, K, bc="dpd"), every=2)
    out = spawn(_work, 4, ("pair" if dims[2] == 4 else "fused", dims, "dpd", N, K, False), backend="nccl")
    for a, b in zip(out[0], base):
        assert a[2] == b[2]


def test_native_mpigpu_two_gpus():
    """One process, two GPUs, peer access, in-kernel halos: `mpigpu -np 2` == 1 GPU trace."""
    import os
    import subprocess
    from mpi_cuda_b200.utils import report
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "bin", "mpigpu")
    one = subprocess.run([exe, "160", "0.001", "20"], capture_output=True, text=True, check=True).stdout
    two = subprocess.run([exe, "-np", "2", "160", "0.001", "20"], capture_output=True, text=True, check=True).stdout
    a = [l for l in one.splitlines() if l.startswith("Step ")]
    b = [l for l in two.splitlines() if l.startswith("Step ")]
    assert len(a) == 10 and a == b          # digit-identical lines, like report.pdf p.15-16
    assert "and 2 processes" in two.splitlines()[0]
