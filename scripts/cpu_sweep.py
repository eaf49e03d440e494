#!/usr/bin/env python
"""Reproduce the reference's CPU sweep tables on this machine (report.pdf p.7-11, 16-17):
OpenMP threads, rank counts, ranks x threads — time, speed-up S = T1/T, efficiency E = S/P —
with the native solvers, and draw the two charts the reference ships as PNGs.

For every N it also prints the phase table of the reference's stage-4 comparison (report.pdf p.16 §4.4:
total / init / compute / boundary / exchange per version).

usage: scripts/cpu_sweep.py [N ...]   (default 128 256 512) -> profiles/cpu_sweeps.md + profiles/*.svg
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mpi_cuda_b200 import WaveProblem                      # noqa: E402
from mpi_cuda_b200.ops import cpu                          # noqa: E402
from mpi_cuda_b200.parallel.decomp import default_dims     # noqa: E402
from mpi_cuda_b200.utils.charts import speedup_efficiency_charts   # noqa: E402
from mpi_cuda_b200.utils.report import speedup_table       # noqa: E402


PHASES = {}      # label -> phases of the fastest repetition (filled by best())


def best(p, reps=3, label=None, **kw):
    runs = []
    for _ in range(reps):
        tr, ph = cpu.solve(p, every=2, **kw)
        runs.append((ph["total"], ph, tr))
    t, ph, tr = min(runs, key=lambda r: r[0])
    if label:
        PHASES[label] = ph
    return t, tr[-1]


def phase_rows(labels):
    out = ["| version | total | init | compute | boundary | exchange |", "|---|---|---|---|---|---|"]
    for lb in labels:
        ph = PHASES[lb]
        out.append("| %s | %.4f | %.4f | %.4f | %.4f | %.4f |" % (lb, ph["total"], ph["init"], ph["compute"],
                                                                  ph.get("boundary", 0.0), ph.get("exchange", 0.0)))
    return "\n".join(out)


def main():
    Ns = [int(a) for a in sys.argv[1:]] or [128, 256, 512]
    ncpu = os.cpu_count() or 1
    threads = [t for t in (1, 2, 4, 8, 16, 32) if t <= ncpu]
    out = ["# CPU sweeps with the native solvers on this box "
           f"({ncpu} hardware threads; the reference used a 20-core POWER8 node)\n",
           "Same tables as report.pdf p.7-11 (time in s, S = T1/T, E = S/P), tau = 0.001, K = 20, L = 1.\n"]
    charts = {}
    for N in Ns:
        p = WaveProblem(N, 1.0, 1e-3, 20)
        PHASES.clear()
        t_seq, last = best(p, threads=1, label="sequential (wave)")
        out.append(f"\n## {N}^3  (sequential {t_seq:.4f} s; step-20 errors {last[2]:.6e} / {last[3]:.6e})\n")
        rows = [("sequential (wave)", 1, t_seq)]
        omp = []
        for t in threads[1:]:
            tt, _ = best(p, threads=t, label=f"OpenMP {t} threads (openmpwave)")
            rows.append((f"OpenMP {t} threads (openmpwave)", t, tt))
            omp.append((t, tt))
        mpi = []
        for r in threads[1:]:
            tt, _ = best(p, threads=1, dims=default_dims(r), label=f"ranks {r} (mpi)")
            rows.append((f"ranks {r} (mpi)", r, tt))
            mpi.append((r, tt))
        hyb = []
        for r in (2, 4):
            t = max(1, ncpu // r)
            if r * t <= ncpu and t > 1:
                tt, _ = best(p, threads=t, dims=default_dims(r), label=f"ranks {r} x {t} threads (mpiomp)")
                rows.append((f"ranks {r} x {t} threads (mpiomp)", r * t, tt))
                hyb.append((r * t, tt))
        out.append(speedup_table(rows))
        out.append(f"\nPhase breakdown at {N}^3 (seconds, max over ranks; same columns as report.pdf p.16 §4.4):\n")
        out.append(phase_rows([r[0] for r in rows]))
        charts[N] = ({"OpenMP": [(1, t_seq)] + omp, "ranks": [(1, t_seq)] + mpi, "ranks x threads": hyb or [(1, t_seq)]}, t_seq)
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    with open(os.path.join(ROOT, "profiles", "cpu_sweeps.md"), "w") as fh:
        fh.write("\n".join(out) + "\n")
    N = Ns[-1]
    sp, ef = speedup_efficiency_charts(charts[N][0], charts[N][1], f"{N}^3, K=20 (native CPU solvers)")
    open(os.path.join(ROOT, "profiles", f"cpu_speedup_{N}.svg"), "w").write(sp)
    open(os.path.join(ROOT, "profiles", f"cpu_efficiency_{N}.svg"), "w").write(ef)
    print("\n".join(out))


if __name__ == "__main__":
    main()
