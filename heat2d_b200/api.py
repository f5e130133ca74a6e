"""The call a user makes: build the right solver for the configuration and the
launch environment, run it, return the result and the final field.

    from heat2d_b200 import HeatConfig, solve
    res, field = solve(HeatConfig(nx=2560, ny=2048, steps=1000))

Under ``torchrun`` (WORLD_SIZE > 1) the same call runs one tile per GPU.
"""
from __future__ import annotations

from typing import Optional, Tuple

from ._lazy import LazyModule

torch = LazyModule("torch")
dist = LazyModule("torch.distributed")

from .config import HeatConfig


def make_solver(cfg: HeatConfig, device=None):
    """Pick the solver class for `cfg.backend` and the current process group."""
    if cfg.backend == "cpu":   # needs neither torch nor a process group
        from .models.cpu import CpuSolver
        return CpuSolver(cfg)
    world = dist.get_world_size() if dist.is_initialized() else 1
    if cfg.backend == "nccl-halo":
        from .parallel.dist import HaloExchangeSolver
        return HaloExchangeSolver(cfg, device)
    if world > 1 and cfg.backend == "fused":
        from .parallel.dist import DistributedHeat2DSolver
        return DistributedHeat2DSolver(cfg, device)
    from .models.heat import Heat2DSolver
    if cfg.backend == "single":
        cfg = cfg.replace(backend="fused", px=1, py=1)
    return Heat2DSolver(cfg, device or "cuda:0")


def solve(cfg: HeatConfig, u0=None, device=None, gather: bool = True) -> Tuple[object, Optional[torch.Tensor]]:
    """Initialise (or load `u0`, a full nx x ny host array), run `cfg.steps`
    steps and return ``(RunResult, final field on the host or None)``.

    A pinned float32 host tensor `u0` on one GPU (fused backend, no convergence
    test) takes the pipelined path: upload, steps and download overlap block by
    block (``Heat2DSolver.solve_pipelined``) and the returned field is pinned."""
    s = make_solver(cfg, device)
    world = dist.get_world_size() if dist.is_initialized() else 1
    if (isinstance(u0, torch.Tensor) and not u0.is_cuda and u0.is_pinned() and u0.dtype == torch.float32
            and world == 1 and cfg.backend in ("fused", "single") and not cfg.conv and cfg.px * cfg.py == 1
            and hasattr(s, "solve_pipelined")):
        try:
            out = torch.empty((cfg.nx, cfg.ny), dtype=torch.float32).pin_memory()
            res = s.solve_pipelined(u0, out, cfg.steps)
            return res, out
        finally:
            s.close()
    try:
        if u0 is None:
            s.initialize()
        else:
            s.load(u0)
        res = s.run()
        field = s.gather() if gather else None
    finally:
        close = getattr(s, "close", None)
        if close:
            close()
    return res, field
