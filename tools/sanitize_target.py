"""Small single-GPU workload for compute-sanitizer (memcheck / racecheck / synccheck):
every kernel family once, on tiny grids, with virtual tiles so the peer-store and flag
code runs too."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from heat2d_b200 import HeatConfig, _native as N
from heat2d_b200.models.heat import Heat2DSolver
from heat2d_b200.ops import kernels as K

for cfg in (HeatConfig(nx=96, ny=160, steps=17, k=8, px=2, py=2, conv=True, interval=5, eps=0.0, chunk_rows=16),
            HeatConfig(nx=70, ny=131, steps=9, k=3, boundary="ghost-ring", arith="strict"),
            HeatConfig(nx=64, ny=64, steps=20, backend="resident"),
            HeatConfig(nx=50, ny=90, steps=5, backend="plain", k=1, conv=True, interval=2, eps=0.0)):
    s = Heat2DSolver(cfg, "cuda:0").initialize()
    s.run()
    ref = K.cpu_initial_grid(cfg.nx, cfg.ny, "exact")
    K.cpu_run(ref, cfg.steps, N.make_phys(cfg.cx, cfg.cy, cfg.boundary, cfg.arith))
    assert np.array_equal(s.gather().numpy(), ref), cfg
    s.close()
torch.cuda.synchronize()
print("sanitize target ok")
