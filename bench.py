#!/usr/bin/env python
"""Headline benchmark: 512^3 wave solve, K time steps, on N B200s of one node.

Metric (BASELINE.json): wall time (s) and GLUP/s at 512^3 K=20, device-timed, max
over ranks.  The reference's published numbers are 0.752 s / 3.55 GLUP/s on one
P100 and 0.505 s / 5.28 GLUP/s on two (readme.md:99-100; report.pdf p.16 §4.4).

A bench "step" is one time step of the scheme (one fused kernel launch per GPU:
stencil + in-kernel P2P halo exchange + error epilogue).  ``--steps K`` runs the
reference's workload exactly when K = 20: Taylor step + 19 leapfrog steps, error
norms on every 2nd step like the reference prints them.  Timed region: K steps,
CUDA events, barrier + synchronize on both sides, max over ranks; repeated
``--reps`` times (each repetition re-initialises the field outside the timed
region) and the median repetition is reported.  The two 1.16 GB field arrays are
far larger than the 126 MB L2, so no flush is needed between iterations.

``--impl reference``: the reference tree ships no source (only readme.md and
report.pdf), so it cannot be installed; prints the "unavailable" line.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PUBLISHED_GLUPS = {1: 3.55, 2: 5.28}      # BASELINE.md §1 (derived from 0.752 s / 0.505 s)
PUBLISHED_WALL_S = {1: 0.752, 2: 0.505}


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU while the timed region runs."""

    def __init__(self, index: int, period: float = 0.005):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.nv = None

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksThrottleReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self) -> dict:
        self._stop_evt.set()
        if self.is_alive():
            self.join(timeout=1.0)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def reference_arm() -> int:
    print(json.dumps({
        "impl": "reference",
        "unavailable": "reference tree has no source code or build (readme.md + report.pdf only): "
                       "pip reports 'neither setup.py nor pyproject.toml found'; published numbers in BASELINE.md stand"}))
    return 0


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=("ours", "reference"))
    ap.add_argument("--N", default="512", help="intervals: one value or NxxNyxNz")
    ap.add_argument("--L", default="1.0", help="box size: one value or LxxLyxLz; 'pi' multiples accepted (e.g. pi, 0.5pi)")
    ap.add_argument("--tau", type=float, default=1e-3)
    ap.add_argument("--bc", default="ddd")
    ap.add_argument("--backend", default="fused", choices=("fused", "nccl"))
    ap.add_argument("--decomp", default="auto", help="auto | slab | block | PxQxR")
    ap.add_argument("--variant", type=int, default=-1)
    ap.add_argument("--zchunk", type=int, default=0)
    ap.add_argument("--reps", type=int, default=7)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--every", type=int, default=2)
    ap.add_argument("--pair", dest="pair", action="store_true", default=True,
                    help="temporal blocking: two time steps per pass over memory (default)")
    ap.add_argument("--no-pair", dest="pair", action="store_false", help="one pass per time step")
    ap.add_argument("--pair-variant", type=int, default=-1)
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm()

    import torch
    import torch.distributed as dist
    from mpi_cuda_b200 import WaveProblem
    from mpi_cuda_b200.models.analytic import oracle_errors
    from mpi_cuda_b200.parallel.decomp import parse_dims
    from mpi_cuda_b200.parallel.launch import init_distributed

    if not torch.cuda.is_available():
        print(json.dumps({"error": "no CUDA device: bench.py needs a B200 (run under gpurun)"}))
        return 1
    rank, world, device = init_distributed("nccl")
    if world != args.gpus:
        if rank == 0:
            print(f"warning: --gpus {args.gpus} but WORLD_SIZE={world}; using {world}", file=sys.stderr)
    K, W = args.steps, max(args.warmup, 3)
    from mpi_cuda_b200.utils.cli import _num
    Ns = [int(v) for v in str(args.N).lower().split("x")]
    Ls = [_num(v) for v in str(args.L).lower().split("x")]
    p = WaveProblem(Ns if len(Ns) == 3 else Ns[0], Ls if len(Ls) == 3 else Ls[0], args.tau, K, bc=args.bc)
    grid_name = "%dx%dx%d" % p.N if len(set(p.N)) > 1 else "%d^3" % p.N[0]
    p.check_cfl()
    dims = parse_dims(args.decomp, world, p.unknowns)

    def barrier():
        torch.cuda.synchronize(device)
        if world > 1:
            dist.barrier()

    def max_over_ranks(x: float) -> float:
        t = torch.tensor([x], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    sampler = ClockSampler(device.index or 0)
    extra = {}
    if args.backend == "fused":
        from mpi_cuda_b200.parallel.fused import FusedSolver, StreamingSession
        s = FusedSolver(p, dims, device, every=args.every, variant=args.variant, zchunk=args.zchunk,
                        use_graph=not args.no_graph, pair=args.pair, pair_variant=args.pair_variant)
        # ---- warm-up: W untimed steps (also loads the module, maps peers) ----
        s.init()
        s._enqueue_steps(1, min(W, K))
        barrier()
        if s.use_graph:
            s.capture()
        times = []
        launches0 = s.kernel_launches
        sampler.start()
        nreps = args.reps
        rep = 0
        while rep < nreps:
            s.init()                       # untimed: layer 0, flags reset, barrier
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            e0.record()
            s.run_steps()                  # EXACTLY K steps
            e1.record()
            barrier()
            times.append(max_over_ranks(e0.elapsed_time(e1)))
            rep += 1
            if rep == 1 and times[0] * nreps < 80.0:
                # a repetition of a few ms gives the clock sampler nothing to see: repeat until the
                # GPU has been under this load for ~80 ms in total (same decision on every rank:
                # times[] holds the max over ranks)
                nreps = int(min(200, max(nreps, 80.0 / max(times[0], 1e-3) + 1)))
        args.reps = nreps
        clocks = sampler.stop()
        launches = (s.kernel_launches - launches0) // args.reps
        s.check_status()
        trace = s.gather_trace()
        extra["plan"] = s.describe()
        # exchange measured inside the kernels during the LAST timed repetition (counters are reset by init):
        # per launch the longest flag wait / halo copy-out over the face CTAs, summed over the launches
        xs = s.exchange_stats()
        extra["exchange"] = {"flag_wait_ms_per_step": xs["flag_wait_s"] * 1e3 / K, "halo_copy_ms_per_step": xs["halo_copy_s"] * 1e3 / K,
                             "max_flag_wait_us": xs["max_flag_wait_s"] * 1e6, "max_halo_copy_us": xs["max_halo_copy_s"] * 1e6,
                             "handshakes": xs["launches"], "measured": "%globaltimer stamps in the face CTAs, max over CTAs and ranks"}
        # ---- end to end through the public step API: per step H2D of the step's
        #      input (time factor, pinned host) and D2H of the step's result ----
        e2e = None
        if not args.no_e2e:
            sess = StreamingSession(s)
            e2e_times = []
            for _ in range(max(3, args.reps // 2)):
                barrier()
                t0 = time.perf_counter()
                sess.start()               # tables H2D from pinned memory + init kernel + barrier
                lag = 2 if s.pair else 1       # keep one launch in flight while earlier results are read
                for n in range(1, K + 1):
                    sess.step_async(n)
                    if n > lag:
                        sess.result(n - lag)
                for n in range(max(1, K - lag + 1), K + 1):
                    sess.result(n)
                barrier()
                e2e_times.append(max_over_ranks(time.perf_counter() - t0))
            sess.close()
            t_e2e = statistics.median(e2e_times)
            e2e = {"value": p.updates / t_e2e / 1e9, "unit": "GLUP/s", "wall_s": t_e2e,
                   "h2d_bytes_per_step": sess.H2D_BYTES_PER_STEP + sess.table_bytes / K,
                   "d2h_bytes_per_step": sess.D2H_BYTES_PER_STEP,
                   "includes": "analytic tables H2D (pinned) + layer-0 init + K x (8 B H2D of the step input, fused step, "
                               "16 B D2H of the step result, host waits for every result; one step kept in flight)"}
    else:
        from mpi_cuda_b200.parallel.exchange import CudaBlockOps, ExchangeSolver
        s = ExchangeSolver(p, dims, lambda pp, b: CudaBlockOps(pp, b, device, args.variant, args.zchunk),
                           every=args.every, backend_name="nccl")
        s.run(time_phases=False)
        times = []
        sampler.start()
        l0 = s.ops.plan.launches
        for _ in range(args.reps):
            r = s.run(time_phases=False)
            times.append(r.phases["loop"] * 1e3)
        clocks = sampler.stop()
        launches = (s.ops.plan.launches - l0) // args.reps
        trace = r.trace
        e2e = None

    t_ms = statistics.median(times)
    glups = p.updates / (t_ms / 1e3) / 1e9
    # correctness guard: the error at the last reporting step must match the closed form
    valid = None
    if trace:
        n, _, mx, l2 = trace[-1]
        omx, ol2 = oracle_errors(p, n)
        valid = bool(abs(mx - omx) <= 1e-13 + 1e-5 * omx and abs(l2 - ol2) <= 1e-13 + 1e-5 * ol2)
    pub = PUBLISHED_GLUPS.get(world, PUBLISHED_GLUPS[2])
    from mpi_cuda_b200.parallel.decomp import plan_all
    from mpi_cuda_b200.utils import roofline
    hbm, peaks = roofline.measured_peaks(ROOT)
    pair_mode = bool(args.pair) and args.backend == "fused"
    rf = roofline.step_roofline(plan_all(p, dims), pair_mode, hbm)
    bytes_per_node, nodes_per_gpu = rf["bytes_per_node_per_step"], rf["nodes_per_gpu"]
    bytes_per_gpu, halo_bytes, t_roof_ms = rf["bytes_per_gpu_per_step"], rf["halo_bytes_per_gpu_per_step"], rf["roofline_ms_per_step"]
    out = {
        "metric": "GLUP/s at %s grid, K=%d steps, fp64, device-timed max over ranks (wall_s = total)" % (grid_name, K),
        "value": glups, "unit": "GLUP/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": t_ms / K, "wall_s": t_ms / 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": glups / pub, "dtype": "fp64", "data": "synthetic (analytic-solution initial layers)",
        "config": {"model": "wave3d leapfrog 7-point", "grid": grid_name, "global_batch": p.n_unknowns,
                   "seq_len": K, "parallelism": f"spatial {dims[0]}x{dims[1]}x{dims[2]}", "backend": args.backend,
                   "bc": args.bc, "L": list(p.L), "temporal_blocking": bool(args.pair), "tau": args.tau, "report_every": args.every,
                   "l2_policy": "inputs larger than L2 (%d x %.2f GB fields per GPU vs 126 MB L2), no flush"
                                % (4 if pair_mode else 2, nodes_per_gpu * 8 / 1e9),
                   "reps": args.reps, "rep_ms": [round(t, 4) for t in times],
                   "cuda_graph": (not args.no_graph) and args.backend == "fused",
                   "published_baseline": {"glups": pub, "wall_s": PUBLISHED_WALL_S.get(world, 0.505),
                                          "hardware": "%d x P100" % min(world, 2)}},
        "roofline": {"hbm_gbs_measured": hbm, "bytes_per_gpu_per_step": bytes_per_gpu,
                     "halo_bytes_per_gpu_per_step": halo_bytes, "roofline_ms_per_step": t_roof_ms,
                     "fraction_of_roofline": t_roof_ms / (t_ms / K), "of": "measured" if peaks else "fallback",
                     "bytes_per_node_per_step": bytes_per_node,
                     "vs_single_pass_roofline": rf["single_pass_roofline_ms_per_step"] / (t_ms / K)},
        "clocks": clocks, "gpu_launches": launches, "validated_vs_oracle": valid,
        "last_error_line": trace[-1] if trace else None,
    }
    if e2e is not None:
        out["e2e"] = e2e
    out.update(extra)
    if rank == 0:
        print(json.dumps(out))
    if world > 1 and dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
