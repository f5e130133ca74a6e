"""heat2d-b200: a Blackwell-native 2-D heat-diffusion framework.

Public surface::

    from heat2d_b200 import HeatConfig, Heat2DSolver, solve
    res, field = solve(HeatConfig(nx=2560, ny=2048, steps=1000, k=8))

Sub-packages: ``models`` (solvers), ``ops`` (kernel bindings), ``parallel``
(decomposition, peer memory, multi-GPU solvers), ``utils`` (dumps, timing, clocks).
"""
from .config import HeatConfig, ConfigError, Launch, best_k, make_schedule

__version__ = "0.1.0"
__all__ = ["HeatConfig", "ConfigError", "Launch", "best_k", "make_schedule", "Heat2DSolver", "CpuSolver", "solve",
           "__version__"]


def __getattr__(name):
    # torch-dependent pieces are imported lazily so `import heat2d_b200` stays cheap
    if name == "Heat2DSolver":
        from .models.heat import Heat2DSolver
        return Heat2DSolver
    if name == "CpuSolver":
        from .models.cpu import CpuSolver
        return CpuSolver
    if name == "solve":
        from .api import solve
        return solve
    raise AttributeError(name)
