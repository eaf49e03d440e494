"""Decomposition planner + the gloo multi-rank backend ("multi-GPU without GPUs").

Reference parity: decomposition invariance — the reference's 1- and 2-GPU traces
are digit-identical (report.pdf p.15-16 §4.3.1-4.3.2).
"""
import pytest

from mpi_cuda_b200 import WaveProblem
from mpi_cuda_b200.ops import cpu
from mpi_cuda_b200.parallel import decomp
from mpi_cuda_b200.parallel.launch import spawn


def test_default_dims():
    assert decomp.default_dims(1) == (1, 1, 1)
    assert decomp.default_dims(2) == (1, 1, 2)
    assert decomp.default_dims(4) == (1, 2, 2)
    assert decomp.default_dims(8) == (1, 2, 4)
    assert decomp.default_dims(8, "slab") == (1, 1, 8)
    assert decomp.parse_dims("2x2x2", 8) == (2, 2, 2)
    with pytest.raises(ValueError):
        decomp.parse_dims("2x2x2", 4)


@pytest.mark.parametrize("N", [512, 1024, 2048])
@pytest.mark.parametrize("dims", [(1, 1, 2), (1, 2, 4), (2, 2, 2), (1, 1, 8)])
def test_uneven_split_covers_everything(N, dims):
    p = WaveProblem(N, 4.0, 1e-3, 1)
    blocks = decomp.plan_all(p, dims)
    assert sum(b.n_local for b in blocks) == p.n_unknowns
    for d in range(3):
        sizes = sorted({b.n[d] for b in blocks})
        assert sizes[-1] - sizes[0] <= 1
    for b in blocks:
        for d in range(3):
            for s in range(2):
                r = b.nbr[d][s]
                if r >= 0:
                    assert blocks[r].nbr[d][1 - s] == b.rank
                    assert blocks[r].n[(d + 1) % 3] == b.n[(d + 1) % 3]
                else:
                    assert b.coord[d] == (0 if s == 0 else dims[d] - 1)


def test_periodic_neighbours_wrap():
    p = WaveProblem(32, 1.0, 1e-3, 1, bc="ddp")
    b0 = decomp.plan_rank(p, (1, 1, 4), 0)
    assert b0.nbr[2] == (3, 1) and b0.nbr[0] == (-1, -1)
    one = decomp.plan_rank(p, (1, 1, 1), 0)
    assert one.nbr[2] == (0, 0) and one.n == (31, 31, 32) and one.lo == (1, 1, 0)
    tz = decomp.local_tables(p, b0)[2]
    assert len(tz) == b0.n[2] + 2 and tz[1] == 0.0        # local index 1 is global node 0
    assert tz[0] == pytest.approx(-tz[2])                  # ghost wraps to node N-1


def test_shell_boxes_partition():
    p = WaveProblem(20, 1.0, 1e-3, 1, bc="ddp")
    for dims in [(2, 2, 2), (1, 2, 2), (1, 1, 1)]:
        for b in decomp.plan_all(p, dims):
            seen = set()
            lo, hi = decomp.interior_box(b)
            boxes = [(lo, hi)] + decomp.shell_boxes(b)
            for (l, h) in boxes:
                for k in range(l[2], h[2]):
                    for j in range(l[1], h[1]):
                        for i in range(l[0], h[0]):
                            assert (i, j, k) not in seen
                            seen.add((i, j, k))
            assert len(seen) == b.n_local


def _gloo_work(rank, world, dims, bc, N, K):
    from mpi_cuda_b200.parallel.exchange import CpuBlockOps, ExchangeSolver
    p = WaveProblem(N, 1.0, 1e-3, K, bc=bc)
    s = ExchangeSolver(p, dims, lambda pp, b: CpuBlockOps(pp, b, 1), every=2, split_shell=True, backend_name="gloo")
    res = s.run()
    return res.trace, res.phases


@pytest.mark.parametrize("world,dims,bc", [(2, (1, 1, 2), "ddd"), (4, (1, 2, 2), "ddp"), (2, (1, 1, 2), "ddp"),
                                           (8, (2, 2, 2), "dpd")])
def test_gloo_rank_count_invariance(world, dims, bc):
    N, K = 28, 6
    base, _ = cpu.solve(WaveProblem(N, 1.0, 1e-3, K, bc=bc), every=2)
    out = spawn(_gloo_work, world, (dims, bc, N, K), backend="gloo")
    for trace, phases in out:
        assert [r[0] for r in trace] == [r[0] for r in base]
        for a, b in zip(trace, base):
            assert a[2] == b[2]
            assert abs(a[3] - b[3]) <= 1e-13 * b[3]
        assert phases["exchange"] > 0


def test_auto_decomposition_policy():
    assert decomp.parse_dims("auto", 8, (511, 511, 511)) == (1, 1, 8)      # 63 planes per rank: slab
    assert decomp.parse_dims("auto", 8, (1023, 1023, 1023)) == (1, 1, 8)   # z faces are the cheap ones
    assert decomp.parse_dims("auto", 8, (2047, 2047, 2047)) == (1, 1, 8)
    assert decomp.parse_dims("auto", 8, (255, 255, 255)) == (1, 1, 8)
    assert decomp.parse_dims("auto", 8, (63, 63, 63)) == (1, 2, 4)         # slabs would be thinner than 8 planes
    assert decomp.parse_dims("auto", 1, (511, 511, 511)) == (1, 1, 1)
    assert decomp.parse_dims("auto", 4) == (1, 2, 2)


def test_local_tables_with_two_ghost_layers():
    """Tables for the wide-halo layout: n + 4 entries, wrapped on periodic axes, 0 beyond a Dirichlet boundary."""
    import numpy as np
    from mpi_cuda_b200.models.analytic import axis_table
    p = WaveProblem(32, 1.0, 1e-3, 1, bc="ddp")
    dims = (1, 1, 4)
    for r in range(4):
        b = decomp.plan_rank(p, dims, r)
        one = decomp.local_tables(p, b, ghosts=1)
        two = decomp.local_tables(p, b, ghosts=2)
        for d in range(3):
            assert len(two[d]) == b.n[d] + 4
            assert np.array_equal(two[d][1:-1], one[d])          # the inner n + 2 entries are the classic table
        full_z = axis_table(p, 2)
        g = np.arange(b.lo[2] - 2, b.lo[2] + b.n[2] + 2) % p.N[2]
        assert np.array_equal(two[2], full_z[g])                 # periodic axis: wrapped
        assert two[0][0] == 0.0 and two[0][-1] == 0.0            # Dirichlet axis: nothing beyond the boundary node
        assert two[0][1] == 0.0 and two[0][-2] == 0.0            # (the boundary node itself)


def test_diagonal_neighbours_are_neighbours_of_neighbours():
    """The wide halo on 3-D blocks sends edge lines to the rank diagonally across an edge; FusedSolver and
    mpigpu find it as the neighbour of the neighbour.  Check that this is symmetric and order-independent."""
    for bc, dims in (("ddd", (2, 2, 2)), ("ppp", (1, 2, 4)), ("dpd", (2, 2, 1)), ("ppp", (1, 1, 1))):
        p = WaveProblem(24, 1.0, 1e-3, 1, bc=bc)
        blocks = decomp.plan_all(p, dims)
        for b in blocks:
            for a, c in ((0, 1), (0, 2), (1, 2)):
                for sa in range(2):
                    for sc in range(2):
                        if b.nbr[a][sa] < 0 or b.nbr[c][sc] < 0:
                            continue
                        d1 = blocks[b.nbr[a][sa]].nbr[c][sc]
                        d2 = blocks[b.nbr[c][sc]].nbr[a][sa]
                        assert d1 == d2 >= 0                                   # either way round
                        back = blocks[blocks[d1].nbr[a][1 - sa]].nbr[c][1 - sc]
                        assert back == b.rank                                  # and the diagonal's diagonal is me
