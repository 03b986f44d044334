"""Property tests of the block decomposition (hypothesis): boxes tile the grid, neighbours are mutual,
update/error ranges and seam planes follow the rules for any N and process grid."""
from hypothesis import given, settings, strategies as st

from wave3d_b200.parallel.decomp import Decomposition

dims_st = st.tuples(st.integers(1, 4), st.integers(1, 4), st.integers(1, 4))


@settings(max_examples=60, deadline=None)
@given(N=st.integers(8, 90), dims=dims_st)
def test_boxes_tile_and_neighbours_are_mutual(N, dims):
    nr = dims[0] * dims[1] * dims[2]
    if any((N + 1) // d < 2 for d in dims):
        return
    dec = Decomposition.create(N, nr, dims)
    boxes = [dec.box(r) for r in range(nr)]
    assert sum(b.n[0] * b.n[1] * b.n[2] for b in boxes) == (N + 1) ** 3
    upd = err = 0
    for b in boxes:
        for a in range(3):
            assert 0 <= b.o[a] and b.o[a] + b.n[a] <= N + 1
            last = b.coords[a] == dims[a] - 1
            assert (b.o[a] + b.n[a] == N + 1) == last
        for f in range(6):
            q = b.nbr[f]
            if q >= 0:
                assert boxes[q].nbr[f ^ 1] == b.rank  # my neighbour's opposite face points back at me
                a = f // 2
                for o in range(3):
                    if o != a:
                        assert boxes[q].n[o] == b.n[o] and boxes[q].o[o] == b.o[o]  # faces match exactly
        upd += (b.hi[0] - b.lo[0] + 1) * (b.hi[1] - b.lo[1] + 1) * (b.hi[2] - b.lo[2] + 1)
        err += max(0, b.ehi[0] - b.elo[0] + 1) * (b.ehi[1] - b.elo[1] + 1) * (b.ehi[2] - b.elo[2] + 1)
        # seam rule: global planes 1 and N-1 cross the periodic boundary
        if b.coords[0] == 0:
            assert b.o[0] + b.send_x[0] == 1
        if b.coords[0] == dims[0] - 1:
            assert b.o[0] + b.send_x[1] == N - 1
    assert upd == (N + 1) * (N - 1) ** 2   # all x planes, interior in y and z
    assert err == (N - 1) ** 3             # global interior
