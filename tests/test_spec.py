"""Spec layer: grid, analytic solution, oracle, CFL, report format, CLI (CPU only).

Mirrors the reference's validation methodology: analytic check,
published delta values, golden trace format.
"""
import math

import numpy as np
import pytest

from mpi_cuda_b200 import WaveProblem, CFLError
from mpi_cuda_b200.models import analytic
from mpi_cuda_b200.utils import report
from mpi_cuda_b200.utils.cli import interpret_positionals, parse

PI = math.pi

# published delta (report.pdf p.7-11) at K=20, tau=1e-3
PUBLISHED = [(128, 1.0, 1.00e-7), (256, 1.0, 2.12e-8), (512, 1.0, 1.40e-9),
             (128, PI, 1.07e-8), (256, PI, 2.62e-9), (512, PI, 6.14e-10)]


@pytest.mark.parametrize("N,L,delta", PUBLISHED)
def test_oracle_matches_published_delta(N, L, delta):
    p = WaveProblem(N, L, 1e-3, 20)
    _, l2 = analytic.oracle_errors(p, 20)
    assert abs(l2 - delta) / delta < 0.01     # published values have 3 digits


def test_oracle_matches_golden_trace():
    p = WaveProblem(512, 1.0, 1e-3, 20)
    got = analytic.oracle_trace(p, every=2)
    assert len(got) == len(report.GOLDEN_512)
    for (n, t, mx, l2), (gn, gt, gmx, gl2) in zip(got, report.GOLDEN_512):
        assert n == gn and abs(t - gt) < 1e-12
        assert abs(mx - gmx) <= 5e-15
        assert abs(l2 - gl2) / gl2 < 2e-6


def test_oracle_high_precision_crosscheck():
    mp = pytest.importorskip("mpmath")
    mp.mp.dps = 60
    p = WaveProblem(512, 1.0, 1e-3, 20)
    h = mp.mpf(1) / 512
    lam = 3 * (4 / h ** 2) * mp.sin(mp.pi * h / 2) ** 2
    tau = mp.mpf("0.001")
    c0, c1 = mp.mpf(1), 1 - tau ** 2 * lam / 2
    for _ in range(2, 21):
        c0, c1 = c1, (2 - tau ** 2 * lam) * c1 - c0
    exact = abs(c1 - mp.cos(mp.pi * mp.sqrt(3) * tau * 20))
    assert abs(analytic.oracle_amplitude_error(p, 20) - float(exact)) / float(exact) < 1e-8


def test_grid_conventions():
    p = WaveProblem(512, 1.0, 1e-3, 20)
    assert p.n_nodes == 513 ** 3 and p.n_unknowns == 511 ** 3
    assert p.updates == 511 ** 3 * 20
    assert abs(p.courant - 0.8868) < 1e-3
    q = WaveProblem(64, 2.0, 1e-3, 10, bc="dpp")
    assert q.unknowns == (63, 64, 64) and q.first_unknown == (1, 0, 0) and q.modes == (1, 2, 2)
    assert abs(q.a_t - PI * math.sqrt(1 / 4 + 4 / 4 + 4 / 4)) < 1e-14


def test_cfl_guard():
    with pytest.raises(CFLError):
        WaveProblem(1024, 1.0, 1e-3, 100).check_cfl()
    assert WaveProblem(1024, PI, 1e-3, 100).check_cfl() < 1.0
    with pytest.raises(CFLError):
        WaveProblem(2048, PI, 1e-3, 1000).check_cfl()
    assert WaveProblem(2048, 4.0, 1e-3, 1000).check_cfl() < 1.0


def test_tables_exact_zero_and_symmetry():
    p = WaveProblem(128, 1.0, 1e-3, 20, bc="ddp")
    sx, sy, sz = analytic.axis_tables(p)
    assert sx[0] == 0.0 and sx[-1] == 0.0
    assert sz[0] == 0.0 and sz[64] == 0.0 and sz[128] == 0.0
    assert abs(sx[64] - 1.0) < 1e-15


def test_step_line_format_is_byte_exact():
    line = report.format_step(20, 0.02, 3.960129e-09, 1.404229e-09)
    assert line == "Step 20, t = 0.020000, Max Error = 3.960129e-09, L2 Error = 1.404229e-09"
    assert report.parse_step(line) == (20, 0.02, 3.960129e-09, 1.404229e-09)
    assert report.banner(512) == "Running test with grid size 512^3"
    assert report.banner(512, 2) == "Running test with grid size 512^3 and 2 processes"


def test_speedup_table_matches_reference_definition():
    # readme.md:88,99,100 — S = T_seq/T, E = S/P
    txt = report.speedup_table([("Sequential", 1, 21.98), ("MPI+CUDA 1 GPU", 1, 0.752), ("MPI+CUDA 2 GPU", 2, 0.505)])
    assert "29.23" in txt and "43.52" in txt and "21.76" in txt


def test_cli_both_spellings():
    assert interpret_positionals(["512", "0.001", "20"]) == dict(N=512, tau=0.001, K=20, L=None, threads=None)
    assert interpret_positionals(["512", "0.001", "20", "1"])["threads"] == 1
    f = interpret_positionals(["512", "1", "0.02", "20"])
    assert f["L"] == 1.0 and f["K"] == 20 and abs(f["tau"] - 1e-3) < 1e-18
    f = interpret_positionals(["1024", "pi", "0.1", "100"])
    assert abs(f["L"] - PI) < 1e-15 and abs(f["tau"] - 1e-3) < 1e-18
    p, args = parse(["128", "0.001", "20", "4", "--bc", "ddp", "--L", "pi"])
    assert p.N == (128,) * 3 and args.threads == 4 and p.bc == "ddp" and abs(p.L[0] - PI) < 1e-15


def test_charts_are_valid_svg():
    import xml.etree.ElementTree as ET
    from mpi_cuda_b200.utils.charts import speedup_efficiency_charts
    # the reference's own headline rows (readme.md:88-100)
    rows = {"MPI": [(1, 39.21), (10, 6.40), (20, 2.73)], "MPI+CUDA": [(1, 0.752), (2, 0.505)]}
    sp, ef = speedup_efficiency_charts(rows, 21.98, "512^3")
    for doc in (sp, ef):
        root = ET.fromstring(doc)
        assert root.tag.endswith("svg")
    assert "MPI+CUDA" in sp and "efficiency" in ef


def test_checkpoint_roundtrip_cpu(tmp_path):
    import torch
    from mpi_cuda_b200.utils import checkpoint
    f = [torch.arange(24, dtype=torch.float64).reshape(2, 3, 4), torch.ones(2, 3, 4, dtype=torch.float64)]
    out = checkpoint.save_state(str(tmp_path / "ck"), 3, 7, f, {"N": (4, 4, 4)})
    assert out.endswith(".rank3.pt")
    st = checkpoint.load_state(str(tmp_path / "ck"), 3)
    assert st["level"] == 7 and torch.equal(st["fields"][0], f[0]) and st["meta"]["N"] == (4, 4, 4)


def test_cli_parses_session_flags():
    p, args = parse(["256", "0.001", "20", "--checkpoint", "/tmp/x", "--checkpoint-at", "10", "--bc", "ddp"])
    assert args.checkpoint == "/tmp/x" and args.checkpoint_at == 10 and p.bc == "ddp"
    p, args = parse(["1024", "pi", "0.1", "100", "--decomp", "1x2x4", "--backend", "fused"])
    assert p.K == 100 and abs(p.tau - 1e-3) < 1e-15 and args.decomp == "1x2x4"


def test_phase_timer_cpu():
    import time
    from mpi_cuda_b200.utils.timing import PhaseTimer, write_chrome_trace
    t = PhaseTimer()
    with t.phase("a"):
        time.sleep(0.01)
    with t.phase("a"):
        time.sleep(0.01)
    with t.phase("b"):
        pass
    r = t.result()
    assert r["a"] >= 0.02 and r["b"] >= 0.0 and set(r) == {"a", "b"}


def test_roofline_arithmetic():
    from mpi_cuda_b200.parallel.decomp import plan_all
    from mpi_cuda_b200.utils import roofline
    p = WaveProblem(512, 1.0, 1e-3, 20)
    one = roofline.step_roofline(plan_all(p, (1, 1, 1)), False, 6539.9)
    assert one["nodes_per_gpu"] == 513 ** 3 and one["halo_bytes_per_gpu_per_step"] == 0
    assert abs(one["roofline_ms_per_step"] - 0.4954) < 1e-3          # 0.495 ms
    two = roofline.step_roofline(plan_all(p, (1, 1, 1)), True, 6539.9)
    assert abs(two["roofline_ms_per_step"] * 3 - one["roofline_ms_per_step"] * 2) < 1e-9
    slab = roofline.step_roofline(plan_all(p, (1, 1, 8)), True, 6539.9)
    assert slab["halo_bytes_per_gpu_per_step"] == 2 * 511 * 511 * 8
    assert slab["roofline_ms_per_step"] < one["roofline_ms_per_step"] / 8 * 1.1


def test_bench_validates_every_line_against_the_published_trace():
    """bench.py's own guard: the headline config is checked line by line against report.pdf p.15-16."""
    import importlib.util
    import os
    from mpi_cuda_b200 import WaveProblem
    from mpi_cuda_b200.utils import report
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(root, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    p = WaveProblem(512, 1.0, 1e-3, 20)
    good = [tuple(g) for g in report.GOLDEN_512]
    ok, how = bench.validate_trace(p, good, 2)
    assert ok and "published" in how
    bad = list(good)
    n, t, mx, l2 = bad[4]
    bad[4] = (n, t, mx + 2e-14, l2)                   # a stale ghost moves Max by more than round-off
    assert bench.validate_trace(p, bad, 2)[0] is False
    assert bench.validate_trace(p, good[:-1], 2)[0] is False        # a missing line is a failure too
    # any other configuration: closed-form oracle on every line
    from mpi_cuda_b200.models.analytic import oracle_trace
    q = WaveProblem(64, 2.0, 1e-3, 10, bc="ddp")
    tr = oracle_trace(q, 2)
    ok, how = bench.validate_trace(q, tr, 2)
    assert ok and "oracle" in how
    tr[2] = (tr[2][0], tr[2][1], tr[2][2] * 1.001, tr[2][3])
    assert bench.validate_trace(q, tr, 2)[0] is False
