"""CPU solver: the oracle behind every GPU test and the in-process stand-in for the
reference's three CPU programs (``mpi_heat2Dn.c``, ``grad1612_mpi_heat.c``,
``grad1612_hybrid_heat.c``).  The decomposition never changes a single bit of the
result (every point is updated from the same five neighbours), so one serial /
OpenMP sweep over the whole grid reproduces what P ranks compute.
"""
from __future__ import annotations

import time
from typing import Optional

import numpy as np

from .. import _native as N
from ..config import HeatConfig
from ..ops import kernels as K


class CpuResult:
    def __init__(self, steps_done, iterations, converged, last_norm, elapsed_s, nx, ny, steps_run=None):
        self.steps_done, self.iterations, self.converged = steps_done, iterations, converged
        self.steps_run = steps_done if steps_run is None else steps_run   # updates applied by this call
        self.last_norm, self.elapsed_s, self.nx, self.ny = last_norm, elapsed_s, nx, ny
        self.launches, self.checks = 0, 0

    @property
    def glups(self) -> float:
        return self.nx * self.ny * self.steps_run / self.elapsed_s / 1e9 if self.elapsed_s > 0 else float("nan")


class CpuSolver:
    def __init__(self, cfg: HeatConfig):
        self.cfg = cfg.validate()
        self.phys = N.make_phys(cfg.cx_eff, cfg.cy_eff, cfg.boundary, cfg.arith)
        self.u: Optional[np.ndarray] = None
        self.steps_done = cfg.start_step

    def initialize(self, mode: Optional[str] = None):
        self.u = K.cpu_initial_grid(self.cfg.nx, self.cfg.ny, mode or self.cfg.init)
        return self

    def load(self, grid: np.ndarray):
        self.u = np.ascontiguousarray(grid, dtype=np.float32).copy()
        return self

    def run(self, steps: Optional[int] = None) -> CpuResult:
        cfg = self.cfg
        steps = cfg.steps if steps is None else steps
        t0 = time.perf_counter()
        done, last = K.cpu_run(self.u, steps, self.phys, cfg.norm if cfg.conv else "none", cfg.interval,
                               cfg.eps, cfg.threads)
        el = time.perf_counter() - t0
        conv = cfg.conv and done < steps
        self.steps_done += done
        return CpuResult(self.steps_done, self.steps_done, conv, last, el, cfg.nx, cfg.ny, done)

    def gather(self, as_numpy: bool = False):
        """The current field on the host: a torch tensor sharing the memory (like the GPU
        solvers' `gather`), or the numpy array itself with ``as_numpy=True`` (no torch import)."""
        if as_numpy:
            return self.u
        import torch
        return torch.from_numpy(self.u)
