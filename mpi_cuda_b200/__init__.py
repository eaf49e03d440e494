"""mpi_cuda_b200 — a Blackwell-native 3-D wave-equation solver framework.

Same capabilities as the AICCer1/MPI-CUDA coursework project (sequential, OpenMP,
MPI, MPI+OpenMP and MPI+CUDA solvers of one explicit leapfrog scheme), re-designed
for 1-8 B200 GPUs: PyTorch host layer, hand-written sm_100a kernels (TMA plane
streaming, in-kernel NVLink P2P halo exchange with flag handshake, fused error
reduction), native C++ CPU solvers as oracle and baseline.
"""
from .models.problem import WaveProblem, CFLError
from .models.solver import solve, BACKENDS
from .utils.report import RunResult, format_step

__version__ = "0.1.0"
__all__ = ["WaveProblem", "CFLError", "solve", "BACKENDS", "RunResult", "format_step"]
