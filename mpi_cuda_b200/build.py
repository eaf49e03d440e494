"""In-tree build of the native components (no JIT cache: the built .so files
travel with the repository snapshot to the GPU box).

Counterpart of the reference's build lines and Makefile (readme.md:34-80).
``python -m mpi_cuda_b200.build [cpu|cuda|tools|all]``.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
NATIVE = Path(__file__).resolve().parent / "_native"
CPU_LIB = NATIVE / "libwave3d_cpu.so"
# W3D_CUDA_LIB: load an alternative build of the CUDA library (kernel experiments: scripts/build_variants.sh)
CUDA_LIB = Path(os.environ["W3D_CUDA_LIB"]).resolve() if os.environ.get("W3D_CUDA_LIB") else NATIVE / "libwave3d_b200.so"


def _run_make(target: str, quiet: bool = True) -> None:
    make = shutil.which("make")
    if make is None:
        raise RuntimeError("make not found")
    env = dict(os.environ)
    env.pop("CXX", None)  # the image's CXX wrapper has no OpenMP runtime; Makefile picks /usr/bin/g++
    proc = subprocess.run([make, "-C", str(ROOT), target], env=env, text=True,
                          stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if proc.returncode != 0:
        raise RuntimeError(f"make {target} failed:\n{proc.stdout[-4000:]}")
    if not quiet:
        print(proc.stdout)


def build_cpu(quiet: bool = True) -> Path:
    _run_make("cpu", quiet)
    return CPU_LIB


def build_cuda(quiet: bool = True) -> Path:
    _run_make("cuda", quiet)
    return CUDA_LIB


def build_tools(quiet: bool = True) -> None:
    _run_make("tools", quiet)


def build_all(quiet: bool = True) -> None:
    _run_make("all", quiet)


def ensure(lib: Path) -> Path:
    """Return the library path, building it on first use if it is missing."""
    if not lib.exists():
        _run_make("cpu" if lib == CPU_LIB else "cuda")
    return lib


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    {"cpu": build_cpu, "cuda": build_cuda, "tools": build_tools, "all": build_all}[what](quiet=False)
