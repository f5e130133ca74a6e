"""Property tests (hypothesis) for the host-side index math: partitioning, neighbour
topology and the launch schedule.  The reference hard-codes these per run in macros and
has no tests at all."""
from hypothesis import given, settings, strategies as st

from heat2d_b200.config import KMAX, make_schedule
from heat2d_b200.parallel.decomp import DIR_DELTA, OPPOSITE, Decomposition, split

settings.register_profile("h2d", max_examples=200, deadline=None)
settings.load_profile("h2d")


@given(st.integers(1, 5000), st.integers(1, 64))
def test_split_partitions_contiguously(n, parts):
    if parts > n:
        return
    blocks = split(n, parts)
    assert len(blocks) == parts
    assert blocks[0][0] == 0 and sum(sz for _, sz in blocks) == n
    for (o0, s0), (o1, _) in zip(blocks, blocks[1:]):
        assert o1 == o0 + s0
    sizes = [sz for _, sz in blocks]
    assert max(sizes) - min(sizes) <= 1
    # the first n % parts blocks carry the extra line (mpi_heat2Dn.c:89-96)
    assert sizes == sorted(sizes, reverse=True)
    assert sizes.count(max(sizes)) == (n % parts or parts)


@given(st.integers(1, 300), st.integers(1, 300), st.integers(1, 6), st.integers(1, 6))
def test_decomposition_topology(nx, ny, px, py):
    if px > nx or py > ny:
        return
    d = Decomposition(nx, ny, px, py)
    assert d.world == px * py
    cells = 0
    for t in d.tiles:
        cells += t.rows * t.cols
        assert d.coords(t.rank) == (t.bx, t.by) and d.rank_of(t.bx, t.by) == t.rank
        # ranks fill down first, then right (grad1612_mpi_heat.c:125-137)
        assert t.rank == t.by * px + t.bx
        for name, (dx, dy) in DIR_DELTA.items():
            q = t.neighbour(name)
            inside = 0 <= t.bx + dx < px and 0 <= t.by + dy < py
            assert (q is not None) == inside
            if q is not None:
                other = d.tile(q)
                assert other.neighbour(OPPOSITE[name]) == t.rank
                assert (other.bx, other.by) == (t.bx + dx, t.by + dy)
                if dx == 1 and dy == 0:
                    assert other.x0 == t.x0 + t.rows and other.y0 == t.y0
                if dx == 0 and dy == 1:
                    assert other.y0 == t.y0 + t.cols and other.x0 == t.x0
        assert d.owner_of(t.x0, t.y0) == t.rank
        assert d.owner_of(t.x0 + t.rows - 1, t.y0 + t.cols - 1) == t.rank
    assert cells == nx * ny


@given(st.integers(1, 100000), st.integers(1, 100000), st.sampled_from([1, 2, 4, 8]))
def test_auto_decomposition_uses_every_gpu(nx, ny, world):
    if nx * ny < world or (nx < world and ny < world):
        return
    try:
        d = Decomposition.auto(nx, ny, world)
    except ValueError:
        # only legitimate when no factorisation fits the grid
        assert all(px > nx or world // px > ny for px in range(1, world + 1) if world % px == 0)
        return
    assert d.world == world
    # no other fitting grid has fewer thin tiles, nor (equally thin) less halo per tile
    def key(px):
        py = world // px
        return ((nx // px < 512) + (ny // py < 512), nx / px + ny / py)
    fits = [px for px in range(1, world + 1) if world % px == 0 and px <= nx and world // px <= ny]
    assert key(d.px) == min(key(px) for px in fits)


@given(st.integers(0, 3000), st.integers(1, KMAX), st.booleans(), st.integers(1, 64), st.integers(0, 500))
def test_schedule_invariants(steps, k, conv, interval, start):
    sched = make_schedule(steps, k, conv, interval, "exact", start)
    assert sum(l.k for l in sched) == steps
    s = start
    for l in sched:
        assert 1 <= l.k <= k
        assert l.first_step == s
        s += l.k
        if l.check:
            assert conv and (s - 1) % interval == 0      # a check launch ENDS on a checked step
    if conv:
        # every step index of the run that is a multiple of `interval` is checked by exactly
        # one launch (grad1612_mpi_heat.c:262-263 tests inside the step loop, step 0 included)
        want = [i for i in range(start, start + steps) if i % interval == 0]
        got = [l.first_step + l.k - 1 for l in sched if l.check]
        assert got == want
        # no launch crosses a checked step
        for l in sched:
            inner = [i for i in range(l.first_step, l.first_step + l.k - 1) if i % interval == 0]
            assert not [i for i in inner if i in want]
    else:
        assert not any(l.check for l in sched)
        # as few launches as k allows, depths differing by at most one (no trailing shallow launch)
        assert len(sched) == -(-steps // k)
        assert not sched or max(l.k for l in sched) - min(l.k for l in sched) <= 1
