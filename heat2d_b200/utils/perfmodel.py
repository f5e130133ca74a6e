"""Analytic cost models.

Two models live here:

* :class:`ClusterModel` — the model the reference report fits to its cluster with mpptest
  (``Report.pdf`` p10-11): per-point compute time ``t_c`` and a message
  cost ``t_s + t_w * L``; used there to argue that 2-D blocks beat 1-D strips
  (``Report.pdf`` p9-14).  Kept so the published tables can be sanity-checked against
  their own model; the constants are the report's.
* :class:`B200Model` — the three-roof model BASELINE.json asks the fused path to be
  reported against: the slowest of fp32 pipe time, HBM time and NVLink halo time, plus a
  per-launch latency floor for small grids.  Denominators are measured values
  (``MEASURED_PEAKS.json`` for HBM; ``profiles/fp32_peak.txt`` and
  ``profiles/p2p_probe_*gpu.json`` for the packed-fp32 rate and the peer store rate).

Nothing here touches a GPU; bench.py and the CLI call :func:`roofline_fraction`.
"""
from __future__ import annotations

import json
import math
import os
from dataclasses import dataclass
from typing import Dict, Optional

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

# Width of one task strip and the overlap rule of csrc/stencil_tk.cu (TK_W, H = roundup4(k)).
STRIP_COLS = 128


def _round_up(v: int, m: int) -> int:
    return (v + m - 1) // m * m


@dataclass(frozen=True)
class ClusterModel:
    """T = steps * (t_c * points_per_rank + messages * (t_s + t_w * words))  [microseconds]."""
    t_c: float = 0.025   # us per grid-point update
    t_s: float = 0.6     # us start-up per message
    t_w: float = 0.9     # us per transferred word (one float)

    def step_us(self, nx: int, ny: int, px: int, py: int) -> float:
        rows, cols = math.ceil(nx / px), math.ceil(ny / py)
        comm = 0.0
        if px > 1:   # a row halo to the north and one to the south (interior rank)
            comm += 2 * (self.t_s + self.t_w * cols)
        if py > 1:   # column halos west and east
            comm += 2 * (self.t_s + self.t_w * rows)
        return self.t_c * rows * cols + comm

    def run_s(self, nx: int, ny: int, px: int, py: int, steps: int) -> float:
        return steps * self.step_us(nx, ny, px, py) * 1e-6

    def speedup(self, nx: int, ny: int, px: int, py: int) -> float:
        return self.step_us(nx, ny, 1, 1) / self.step_us(nx, ny, px, py)


@dataclass(frozen=True)
class B200Model:
    hbm_gbs: float = 6551.0           # MEASURED_PEAKS.json (copy bandwidth, read + write bytes)
    fp32_lane_fma_per_s: float = 34.9e12   # packed FFMA2, whole chip (profiles/fp32_peak.txt)
    nvlink_store_gbs: float = 774.0   # peer stores, one direction, large transfers
    launch_us: float = 4.0            # back-to-back launch cadence of a short kernel
    lane_slots_per_update: float = 5.0     # 2 adds, 1 mul, 2 fma per point and step

    @staticmethod
    def from_measured(path: Optional[str] = None) -> "B200Model":
        path = path or os.path.join(_ROOT, "MEASURED_PEAKS.json")
        try:
            with open(path) as f:
                hbm = float(json.load(f).get("hbm_gbs") or 0.0)
        except (OSError, ValueError):
            hbm = 0.0
        return B200Model(hbm_gbs=hbm) if hbm > 0 else B200Model()

    # ---- per-GPU cost of ONE launch of k steps on a rows x cols tile --------------
    def launch_costs_us(self, rows: int, cols: int, k: int, chunk_rows: int = 0,
                        row_neighbours: int = 0, col_neighbours: int = 0) -> Dict[str, float]:
        k = max(1, k)
        h = _round_up(k, 4)                       # overlap columns per side of a strip
        stored = STRIP_COLS - 2 * h               # columns a strip keeps
        strips = max(1, math.ceil(cols / stored))
        r = chunk_rows if chunk_rows > 0 else rows
        chunks = max(1, math.ceil(rows / r))
        rows_swept = rows + chunks * 2 * k        # every chunk refills its 2k pipeline rows
        cols_swept = strips * STRIP_COLS
        updates = rows * cols * k                 # useful point updates
        # fp32 pipe: every level of every swept cell costs the 5 lane slots
        pipe_s = rows_swept * cols_swept * k * self.lane_slots_per_update / self.fp32_lane_fma_per_s
        # HBM: read the swept window once, write the owned cells once
        bytes_hbm = 4.0 * rows_swept * cols_swept + 4.0 * rows * cols
        hbm_s = bytes_hbm / (self.hbm_gbs * 1e9)
        # NVLink: k rows per row neighbour, k (rounded to full 32-column lines) columns per column neighbour
        bytes_nvl = 4.0 * (row_neighbours * k * cols + col_neighbours * 32 * rows)
        nvl_s = bytes_nvl / (self.nvlink_store_gbs * 1e9)
        return {"pipe_us": pipe_s * 1e6, "hbm_us": hbm_s * 1e6, "nvlink_us": nvl_s * 1e6,
                "launch_us": self.launch_us, "updates": float(updates)}

    def bound_glups(self, rows: int, cols: int, k: int, chunk_rows: int = 0,
                    row_neighbours: int = 0, col_neighbours: int = 0, blocking_overhead: bool = True) -> Dict:
        """Roofline of one GPU: updates / slowest roof.  With ``blocking_overhead=False``
        the redundant strip / pipeline-fill work is left out (the algorithmic roof)."""
        if blocking_overhead:
            c = self.launch_costs_us(rows, cols, k, chunk_rows, row_neighbours, col_neighbours)
        else:
            upd = float(rows) * cols * max(1, k)
            c = {"pipe_us": upd * self.lane_slots_per_update / self.fp32_lane_fma_per_s * 1e6,
                 "hbm_us": 8.0 * rows * cols / (self.hbm_gbs * 1e9) * 1e6,
                 "nvlink_us": 4.0 * (row_neighbours * k * cols + col_neighbours * k * rows)
                 / (self.nvlink_store_gbs * 1e9) * 1e6,
                 "launch_us": self.launch_us, "updates": upd}
        roofs = {n: c[n] for n in ("pipe_us", "hbm_us", "nvlink_us", "launch_us")}
        limiter = max(roofs, key=roofs.get)
        return {"glups": c["updates"] / roofs[limiter] * 1e-3, "limiter": limiter[:-3], **roofs}


def roofline_fraction(glups_total: float, nx: int, ny: int, px: int, py: int, k: int,
                      chunk_rows: int = 0, model: Optional[B200Model] = None) -> Dict:
    """Achieved fraction of the three-roof bound for a px x py run (whole job)."""
    model = model or B200Model.from_measured()
    rows, cols = math.ceil(nx / px), math.ceil(ny / py)
    rn, cn = min(2, px - 1), min(2, py - 1)   # an interior tile of the process grid
    with_ovh = model.bound_glups(rows, cols, k, chunk_rows, rn, cn, True)
    algo = model.bound_glups(rows, cols, k, 0, rn, cn, False)
    world = px * py
    return {
        "bound_glups_with_blocking": with_ovh["glups"] * world, "limiter": with_ovh["limiter"],
        "bound_glups_algorithm": algo["glups"] * world, "limiter_algorithm": algo["limiter"],
        "fraction_with_blocking": glups_total / (with_ovh["glups"] * world) if with_ovh["glups"] else None,
        "fraction_algorithm": glups_total / (algo["glups"] * world) if algo["glups"] else None,
        "k1_hbm_bound_glups": model.hbm_gbs / 8.0 * world,
    }
