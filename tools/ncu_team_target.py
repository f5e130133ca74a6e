"""ncu target: one independent-strip launch and one team launch per cluster size, strict-fma and fp32.

    python tools/ncu_team_target.py [n]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from heat2d_b200 import HeatConfig
from heat2d_b200.models.heat import Heat2DSolver

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
for arith in ("fp32", "strict-fma"):
    s = Heat2DSolver(HeatConfig(nx=n, ny=n, steps=8, k=8, arith=arith, layout="v4"), "cuda:0").initialize()
    p = s.engines[0].plan
    p.launch(0, 8, 1, "none", 0, 0)
    for cl in (1, 2):
        p.launch_team(0, 8, cl, 0)
    torch.cuda.synchronize()
    print(arith, "ok", flush=True)
    s.close()
