"""Domain decomposition and neighbour topology (pure index math, no MPI).

Replaces ``MPI_Cart_create`` / ``MPI_Cart_shift`` and the ``xs[]/ys[]`` tables of
``grad1612_mpi_heat.c:47-52,74-81,125-137`` and the strip partition of
``mpi_heat2Dn.c:89-116``.

Conventions (identical to the reference):

* ``px`` (= GRIDX) splits the rows (x, slow index), ``py`` (= GRIDY) the columns.
* rank ``r`` owns row-block ``r % px`` and column-block ``r // px`` — ranks fill
  *down first, then right* (``ys[r] = (r / GRIDX)·ycell``, ``xs[r] = (r % GRIDX)·xcell``).
* NORTH is the smaller row index, WEST the smaller column index; a missing
  neighbour is ``None`` (``MPI_PROC_NULL``).
* uneven sizes: the first ``n % parts`` blocks get one extra line
  (``mpi_heat2Dn.c:89-96``); the grad1612 programs refuse them (``:60-63``).

On an NVSwitch box every GPU pair has the same bandwidth, so rank placement
(``reorder=1``) is irrelevant and not modelled.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

DIRS = ("N", "S", "W", "E", "NW", "NE", "SW", "SE")
DIR_INDEX = {d: i for i, d in enumerate(DIRS)}
# (row-block delta, column-block delta)
DIR_DELTA = {"N": (-1, 0), "S": (1, 0), "W": (0, -1), "E": (0, 1),
             "NW": (-1, -1), "NE": (-1, 1), "SW": (1, -1), "SE": (1, 1)}
OPPOSITE = {"N": "S", "S": "N", "W": "E", "E": "W", "NW": "SE", "SE": "NW", "NE": "SW", "SW": "NE"}


def split(n: int, parts: int) -> List[Tuple[int, int]]:
    """(start, size) of each of `parts` contiguous blocks of `n` lines."""
    if parts < 1 or n < parts:
        raise ValueError(f"cannot split {n} lines into {parts} blocks")
    base, extra = divmod(n, parts)
    out, off = [], 0
    for i in range(parts):
        sz = base + 1 if i < extra else base
        out.append((off, sz))
        off += sz
    return out


@dataclass(frozen=True)
class TileSpec:
    rank: int
    bx: int          # row-block index   (0 .. px-1)
    by: int          # column-block index (0 .. py-1)
    x0: int
    y0: int
    rows: int
    cols: int
    neighbours: Dict[str, Optional[int]] = field(default_factory=dict, compare=False)

    def neighbour(self, d: str) -> Optional[int]:
        return self.neighbours.get(d)


class Decomposition:
    """px x py block decomposition of an nx x ny grid."""

    def __init__(self, nx: int, ny: int, px: int = 1, py: int = 1, uneven: bool = True):
        if px < 1 or py < 1:
            raise ValueError("px and py must be >= 1")
        if not uneven and (nx % px or ny % py):
            raise ValueError(f"ERROR: ({nx}/{px}) or ({ny}/{py}) is not an integer")
        self.nx, self.ny, self.px, self.py = nx, ny, px, py
        self.row_blocks = split(nx, px)
        self.col_blocks = split(ny, py)
        self.tiles: List[TileSpec] = []
        for r in range(px * py):
            bx, by = self.coords(r)
            x0, rows = self.row_blocks[bx]
            y0, cols = self.col_blocks[by]
            nb = {}
            for d, (dx, dy) in DIR_DELTA.items():
                nb[d] = self.rank_of(bx + dx, by + dy)
            self.tiles.append(TileSpec(r, bx, by, x0, y0, rows, cols, nb))

    @property
    def world(self) -> int:
        return self.px * self.py

    def coords(self, rank: int) -> Tuple[int, int]:
        return rank % self.px, rank // self.px

    def rank_of(self, bx: int, by: int) -> Optional[int]:
        if 0 <= bx < self.px and 0 <= by < self.py:
            return by * self.px + bx
        return None

    def tile(self, rank: int) -> TileSpec:
        return self.tiles[rank]

    def min_extent(self) -> int:
        return min(min(sz for _, sz in self.row_blocks), min(sz for _, sz in self.col_blocks))

    def owner_of(self, x: int, y: int) -> int:
        bx = next(i for i, (o, s) in enumerate(self.row_blocks) if o <= x < o + s)
        by = next(i for i, (o, s) in enumerate(self.col_blocks) if o <= y < o + s)
        return self.rank_of(bx, by)

    @staticmethod
    def strips(nx: int, ny: int, workers: int) -> "Decomposition":
        """The 1-D row-strip layout of the given program (mpi_heat2Dn.c)."""
        return Decomposition(nx, ny, px=workers, py=1, uneven=True)

    @staticmethod
    def auto(nx: int, ny: int, world: int, min_rows: int = 512) -> "Decomposition":
        """Process grid for `world` GPUs of one NVSwitch box: the 2-D block decomposition of
        the reference (GRIDX x GRIDY, grad1612_mpi_heat.c:47-52) with the most square tiles,
        i.e. the least halo volume, preferring to split rows when two grids tie (a row
        band is contiguous in memory).

        With the column bands travelling through packed ghost buffers every layout runs at
        the same speed on B200 (32768^2, fp32: 4x1 16484, 2x2 16478, 1x4 16242 GLUP/s on four
        GPUs; 1x2 = 2x1 = 8583 on two), so the choice is about halo bytes and tile shape
        only.  Tiles keep at least `min_rows` rows and columns where the grid allows it
        (short tiles are all pipeline fill)."""
        best = None
        for px in range(1, world + 1):
            if world % px:
                continue
            py = world // px
            if px > nx or py > ny:
                continue
            thin = (nx // px < min_rows) + (ny // py < min_rows)   # avoid tiles that are too short / narrow
            cost = nx / px + ny / py                               # halo volume per tile
            key = (thin, cost, -px)
            if best is None or key < best[0]:
                best = (key, px, py)
        if best is None:
            raise ValueError(f"no px x py == {world} fits a {nx}x{ny} grid")
        return Decomposition(nx, ny, best[1], best[2])

    def describe(self) -> str:
        lines = [f"{self.px}x{self.py} tiles over {self.nx}x{self.ny}"]
        for t in self.tiles:
            nb = ", ".join(f"{d}={t.neighbours[d]}" for d in ("N", "S", "E", "W"))
            lines.append(f"  rank {t.rank}: rows [{t.x0},{t.x0 + t.rows}) cols [{t.y0},{t.y0 + t.cols}) {nb}")
        return "\n".join(lines)
