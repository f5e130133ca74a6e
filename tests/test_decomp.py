"""Decomposition / topology vs the reference's formulas
(grad1612_mpi_heat.c:47-52,74-81,125-137; mpi_heat2Dn.c:89-116)."""
import pytest

from heat2d_b200.parallel.decomp import Decomposition, OPPOSITE, split


def test_split_matches_given_program():
    # averow = NX / nw; extra = NX % nw; first `extra` workers get +1 (mpi_heat2Dn.c:89-96)
    assert split(640, 3) == [(0, 214), (214, 213), (427, 213)]
    assert split(10, 4) == [(0, 3), (3, 3), (6, 2), (8, 2)]
    with pytest.raises(ValueError):
        split(3, 4)


@pytest.mark.parametrize("nx,ny,px,py", [(10, 10, 2, 2), (640, 512, 4, 4), (2560, 2048, 2, 4), (80, 64, 1, 8)])
def test_origins_match_xs_ys_tables(nx, ny, px, py):
    d = Decomposition(nx, ny, px, py, uneven=False)
    xcell, ycell = nx // px, ny // py
    for r, t in enumerate(d.tiles):
        assert t.x0 == (r % px) * xcell      # xs[r]
        assert t.y0 == (r // px) * ycell     # ys[r]
        assert (t.rows, t.cols) == (xcell, ycell)


def test_neighbours_like_cart_shift():
    d = Decomposition(12, 12, 3, 2)
    # rank = by * px + bx; NORTH/SOUTH move along x (bx -/+ 1), WEST/EAST along y
    t = d.tile(1)  # bx=1, by=0
    assert (t.neighbours["N"], t.neighbours["S"]) == (0, 2)
    assert (t.neighbours["W"], t.neighbours["E"]) == (None, 4)
    assert (t.neighbours["NE"], t.neighbours["SE"], t.neighbours["NW"]) == (3, 5, None)
    for a in d.tiles:
        for dname, nb in a.neighbours.items():
            if nb is not None:
                assert d.tile(nb).neighbours[OPPOSITE[dname]] == a.rank


def test_tiles_partition_the_grid():
    d = Decomposition(37, 53, 3, 4)
    seen = set()
    for t in d.tiles:
        for x in range(t.x0, t.x0 + t.rows):
            for y in range(t.y0, t.y0 + t.cols):
                assert (x, y) not in seen
                seen.add((x, y))
    assert len(seen) == 37 * 53
    assert d.owner_of(36, 52) == d.world - 1


def test_divisibility_rule_of_grad1612():
    with pytest.raises(ValueError, match="is not an integer"):
        Decomposition(10, 10, 3, 2, uneven=False)
    Decomposition(10, 10, 3, 2, uneven=True)


def test_strips_and_auto():
    d = Decomposition.strips(640, 640, 3)
    assert [(t.x0, t.rows) for t in d.tiles] == [(0, 214), (214, 213), (427, 213)]
    assert all(t.cols == 640 for t in d.tiles)
    a = Decomposition.auto(32768, 32768, 8)
    assert (a.px, a.py) == (4, 2)          # most square blocks, rows split first on a tie
    assert (Decomposition.auto(32768, 32768, 4).px, Decomposition.auto(32768, 32768, 4).py) == (2, 2)
    assert (Decomposition.auto(32768, 32768, 2).px, Decomposition.auto(32768, 32768, 2).py) == (2, 1)
    b = Decomposition.auto(640, 65536, 8)  # too few rows to split: column strips
    assert (b.px, b.py) == (1, 8)
    c = Decomposition.auto(65536, 640, 8)
    assert (c.px, c.py) == (8, 1)
    assert Decomposition.auto(100, 100, 1).world == 1
