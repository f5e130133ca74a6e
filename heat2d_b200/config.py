"""Run configuration: the reference's compile-time ``#define`` blocks as runtime options.

Reference: ``grad1612_mpi_heat.c:5-21`` (NXPROB, NYPROB, STEPS, GRIDX, GRIDY,
CONVERGENCE, INTERVAL, SENSITIVITY, CX, CY, DEBUG), ``grad1612_hybrid_heat.c:12-24``
(NUMTHREADS), ``grad1612_cuda_heat.cu:6-13`` (BLOCKX/BLOCKY have no equivalent: the
launch shape is fixed by the kernels), ``mpi_heat2Dn.c:29-44`` (float coefficients).
"""
from __future__ import annotations

import dataclasses
from dataclasses import dataclass
from typing import Optional

from ._native import ARITH, BOUNDARY, INIT, KMAX, NORM

BACKENDS = ("fused", "single", "plain", "nccl-halo", "resident", "cpu")
DUMPS = ("none", "prtdat", "rowmajor", "binary")
CONV_MODES = ("exact", "fused-k")
LAYOUTS = {"auto": 0, "v4": 1, "ab2": 2, "ab4": 3, "v8": 4}


class ConfigError(ValueError):
    pass


def best_k(arith: str, tile_points: int) -> int:
    """Sweep depth (time steps per launch) with the highest measured throughput on B200 for a
    tile of `tile_points` cells per GPU.  fp32: 8 — the kernel needs the 8-fold cut in HBM
    traffic.  Strict arithmetic is bound by instruction dispatch, not by HBM (DRAM 17 % busy at
    k = 8); at k <= 4 the column margin is 4 instead of 8, the pipeline fills in 8 rows instead
    of 16, 20 steps are 5 x 4 instead of 7 + 7 + 6, and the state of 4 levels leaves registers
    for 8 columns per lane (the v8 layout: 8 of 256 columns recomputed).  strict-fma, GLUP/s,
    k = 8 / k = 4 in v4 / k = 4 in v8: 32768^2 1322 / 1366 / 1438; 16384^2 1264 / 1320 / 1322;
    8192 x 16384 (one tile of eight) 1219 / 1220 / 1280; 8192^2 1147 / 1169 / -; 2 x 2 GPUs
    5010 / 5215 / - (profiles/r2/probe_strict-fma_*_ksweep.json, *_v8.json,
    scaling_sweep_n4_ks.jsonl).  k = 3 and k = 2 are slower everywhere."""
    if arith == "fp32":
        return 8
    return 4 if tile_points >= (1 << 27) else 8


@dataclass
class HeatConfig:
    """Everything a run needs.  Defaults are the reference's defaults."""

    nx: int = 10                 # NXPROB: rows (slow index)
    ny: int = 10                 # NYPROB: columns (contiguous index)
    steps: int = 100             # STEPS
    px: int = 1                  # GRIDX: tiles along x (rows)
    py: int = 1                  # GRIDY: tiles along y (columns)
    conv: bool = False           # CONVERGENCE
    interval: int = 20           # INTERVAL
    eps: float = 0.1             # SENSITIVITY
    norm: str = "l2sq"           # l2sq = grad1612_mpi_heat.c:264-269; linf = north-star wording
    conv_mode: str = "exact"     # exact: pre-step field + 0-based index, as the reference
    cx: float = 0.1
    cy: float = 0.1
    cx_float: bool = False       # coefficients rounded to float first (mpi_heat2Dn.c:41-44)
    boundary: str = "fixed"      # fixed | ghost-ring
    arith: str = "fp32"          # fp32 | strict | strict-fma
    init: str = "exact"          # ref-int32 | ref-uint32 | exact | zero
    k: int = 8                   # time steps per launch / per halo exchange
    backend: str = "fused"
    halo: int = 8                # ghost depth of the tile arrays (>= k)
    chunk_rows: int = 0          # 0 = let the plan choose
    uneven: bool = True          # allow non-divisible splits (mpi_heat2Dn.c:89-116 style)
    dump: str = "none"
    out_dir: str = "."
    threads: int = 1             # CPU backend only (NUMTHREADS)
    debug: bool = False
    restart: Optional[str] = None
    start_step: int = 0
    ref_conv_bug: bool = False   # replay the stale-`i` predicate of grad1612_mpi_heat.c:263
    spin_timeout_s: float = 10.0
    dataflow: bool = False       # one dataflow launch per run of equal-k sweeps
                                 # (single tile, no convergence test; h2d_plan_run_dataflow)
    layout: str = "auto"         # lane layout of the kernels (csrc/tk_task.cuh): auto | v4 | ab2 | ab4 (fp32) | v8
                                 # (strict arithmetic with k <= 4; 8 columns per lane); same bits, different
                                 # speed; a layout the arithmetic / depth does not have falls back to v4
    team: str = "off"            # on: sweeps of k >= 5 without neighbours / norm run as cluster launches whose
                                 # strips share their edge columns through shared memory (csrc/tk_team.cuh);
                                 # same bits; faster only in strict arithmetic on very large tiles

    def validate(self, world_size: Optional[int] = None) -> "HeatConfig":
        if self.nx < 1 or self.ny < 1:
            raise ConfigError(f"grid {self.nx}x{self.ny} is empty")
        if self.steps < 0:
            raise ConfigError("steps must be >= 0")
        if self.px < 1 or self.py < 1:
            raise ConfigError("px and py must be >= 1")
        if self.px > self.nx or self.py > self.ny:
            raise ConfigError(f"process grid {self.px}x{self.py} has more tiles than grid lines")
        if world_size is not None and self.backend not in ("single", "resident", "cpu"):
            if self.px * self.py != world_size:
                # grad1612_mpi_heat.c:55-58
                raise ConfigError(
                    f"ERROR: the number of tasks must be equal to {self.px * self.py}.")
        if not self.uneven and (self.nx % self.px or self.ny % self.py):
            # grad1612_mpi_heat.c:60-63
            raise ConfigError(
                f"ERROR: ({self.nx}/{self.px}) or ({self.ny}/{self.py}) is not an integer")
        for name, table in (("boundary", BOUNDARY), ("arith", ARITH), ("init", INIT), ("norm", NORM)):
            if getattr(self, name) not in table:
                raise ConfigError(f"{name} must be one of {sorted(table)}")
        if self.backend not in BACKENDS:
            raise ConfigError(f"backend must be one of {BACKENDS}")
        for dname in self.dump.split(","):
            if dname not in DUMPS:
                raise ConfigError(f"dump must be a comma list of {DUMPS}")
        if self.conv_mode not in CONV_MODES:
            raise ConfigError(f"conv_mode must be one of {CONV_MODES}")
        if self.team not in ("on", "off"):
            raise ConfigError("team must be 'on' or 'off'")
        if self.layout not in LAYOUTS:
            raise ConfigError(f"layout must be one of {sorted(LAYOUTS)}")
        if not 1 <= self.k <= KMAX:
            raise ConfigError(f"k must be in [1, {KMAX}]")
        if self.halo < self.k:
            raise ConfigError("halo must be >= k")
        if self.conv and self.interval < 1:
            raise ConfigError("interval must be >= 1")
        if self.threads < 1:
            raise ConfigError("threads must be >= 1")
        return self

    @property
    def cx_eff(self) -> float:
        """Coefficient as the double the arithmetic uses."""
        import numpy as np
        return float(np.float32(self.cx)) if self.cx_float else float(self.cx)

    @property
    def cy_eff(self) -> float:
        import numpy as np
        return float(np.float32(self.cy)) if self.cx_float else float(self.cy)

    def replace(self, **kw) -> "HeatConfig":
        return dataclasses.replace(self, **kw)

    def to_dict(self) -> dict:
        return dataclasses.asdict(self)


@dataclass(frozen=True)
class Launch:
    """One kernel launch of the schedule: `k` steps; `check` = evaluate the
    convergence norm on its last sub-step."""
    k: int
    check: bool
    first_step: int  # 0-based index of the first step this launch performs


def make_schedule(steps: int, k: int, conv: bool = False, interval: int = 20,
                  conv_mode: str = "exact", start_step: int = 0):
    """Split `steps` time steps into launches of at most `k` steps.

    Convergence is evaluated on 0-based step indices that are multiples of
    `interval` (the intent of ``grad1612_mpi_heat.c:263``): a launch always ENDS on
    a checked step, whose (new - old) norm is reduced inside that launch.
    `conv_mode` does not change the schedule; it decides what the solver returns
    on convergence (``exact``: the pre-step field, as the reference;
    ``fused-k``: the post-step field).
    """
    out = []
    s = start_step
    end = start_step + steps

    def stretch(n: int, check_last: bool):
        # as few launches as k allows, of (nearly) equal depth: 20 steps at k = 8 -> 7 + 7 + 6,
        # not 8 + 8 + 4 (a shallow launch is HBM-bound, a deep one is not)
        nonlocal s
        parts = -(-n // k)
        base, extra = divmod(n, parts)
        for i in range(parts):
            kk = base + (1 if i < extra else 0)
            out.append(Launch(kk, check_last and i == parts - 1, s))
            s += kk

    while s < end:
        if not conv:
            stretch(end - s, False)
            continue
        nxt = ((s + interval - 1) // interval) * interval  # next checked index >= s
        if nxt >= end:
            stretch(end - s, False)
            continue
        stretch(nxt - s + 1, True)  # up to and including the checked step: the last launch ends on it
    return out
