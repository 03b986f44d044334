"""Decomposition arithmetic."""
import pytest

import wave3d_b200 as w
from wave3d_b200.parallel.decomp import Decomposition, dims_create, parse_dims


@pytest.mark.parametrize("n,expect", [(1, (1, 1, 1)), (2, (2, 1, 1)), (3, (3, 1, 1)), (4, (2, 2, 1)), (6, (3, 2, 1)),
                                      (8, (2, 2, 2)), (12, (3, 2, 2)), (16, (4, 2, 2)), (7, (7, 1, 1))])
def test_dims_create_matches_mpi(n, expect):
    assert dims_create(n) == expect  # ref: MPI_Dims_create, cuda_sol.cpp:478


def test_parse_dims():
    assert parse_dims("8x1x1") == (8, 1, 1)
    assert parse_dims("2,2,2") == (2, 2, 2)
    with pytest.raises(ValueError):
        parse_dims("2x2")
    with pytest.raises(ValueError):
        parse_dims("0x1x1")


def test_rank_coords_row_major_last_fastest():
    d = Decomposition.create(63, 8)
    assert d.coords_of(0) == (0, 0, 0) and d.coords_of(1) == (0, 0, 1) and d.coords_of(2) == (0, 1, 0)
    assert d.coords_of(4) == (1, 0, 0)
    for r in range(8):
        assert d.rank_of(d.coords_of(r)) == r


def test_extents_remainder_to_last():
    # N=128 -> 129 points over 2 ranks: 64 + 65, origins 0 and 64 (ref: cuda_sol.cpp:485-489)
    d = Decomposition.create(128, 8)
    b0, b7 = d.box(0), d.box(7)
    assert list(b0.n) == [64, 64, 64] and list(b0.o) == [0, 0, 0]
    assert list(b7.n) == [65, 65, 65] and list(b7.o) == [64, 64, 64]
    d3 = Decomposition.create(32, 3)  # 33 over 3: 11, 11, 11
    assert [d3.box(r).n[0] for r in range(3)] == [11, 11, 11]
    d3 = Decomposition.create(33, 3)  # 34 over 3: 11, 11, 12
    assert [d3.box(r).n[0] for r in range(3)] == [11, 11, 12]
    assert [d3.box(r).o[0] for r in range(3)] == [0, 11, 22]


def test_neighbours_periodic_x_open_yz():
    d = Decomposition.create(63, 8)  # 2x2x2
    b = d.box(0)
    assert b.nbr[0] == d.rank_of((1, 0, 0)) and b.nbr[1] == d.rank_of((1, 0, 0))  # x wraps
    assert b.nbr[2] == -1 and b.nbr[3] == d.rank_of((0, 1, 0))
    assert b.nbr[4] == -1 and b.nbr[5] == d.rank_of((0, 0, 1))
    one = Decomposition.create(63, 1).box(0)
    assert one.nbr[0] == 0 and one.nbr[1] == 0 and list(one.nbr[2:]) == [-1, -1, -1, -1]  # self-wrap


def test_update_and_error_ranges_and_seam_planes():
    d = Decomposition.create(63, 8)
    lo, hi = d.box(0), d.box(7)
    # x: all owned planes are advanced; error set excludes global 0 and N
    assert (lo.lo[0], lo.hi[0]) == (0, 31) and (lo.elo[0], lo.ehi[0]) == (1, 31)
    assert (hi.lo[0], hi.hi[0]) == (0, 31) and (hi.elo[0], hi.ehi[0]) == (0, 30)
    # y/z: global interior only
    assert (lo.lo[1], lo.hi[1]) == (1, 31) and (hi.lo[2], hi.hi[2]) == (0, 30)
    # seam rule (ref: cuda_sol.cpp:232): bottom rank sends global 1 down, top rank sends global N-1 up
    assert list(lo.send_x) == [1, 31] and list(hi.send_x) == [0, 30]
    one = Decomposition.create(63, 1).box(0)
    assert list(one.send_x) == [1, 62]


def test_layout_alignment():
    for nr, dims in [(1, None), (8, None), (4, (1, 2, 2)), (3, (1, 1, 3))]:
        d = Decomposition.create(100, nr, dims)
        for r in range(nr):
            b, lay = d.box(r), d.layout(r)
            assert lay.pitch % 16 == 0 and lay.oz + b.lo[2] == 16  # first updated z cell is 128-byte aligned
            assert lay.pitch >= lay.oz + b.n[2] + 1 and lay.oz >= 1
            assert lay.rows == b.n[1] + 2 and lay.planes == b.n[0] + 2
            assert lay.elems == lay.planes * lay.rows * lay.pitch


def test_too_many_ranks_rejected():
    with pytest.raises(ValueError):
        Decomposition.create(4, 8, (8, 1, 1)).box(0)


def test_boxes_tile_the_grid_exactly():
    N = 37
    for nr, dims in [(6, None), (8, (8, 1, 1)), (8, (1, 2, 4)), (12, None)]:
        d = Decomposition.create(N, nr, dims)
        seen = set()
        for r in range(nr):
            b = d.box(r)
            for a in range(3):
                assert b.o[a] + b.n[a] <= N + 1
            cells = b.n[0] * b.n[1] * b.n[2]
            key = (b.o[0], b.o[1], b.o[2])
            assert key not in seen
            seen.add(key)
        tot = sum(d.box(r).n[0] * d.box(r).n[1] * d.box(r).n[2] for r in range(nr))
        assert tot == (N + 1) ** 3
