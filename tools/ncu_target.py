"""Small program for ncu captures: one launch each of several kernel variants.

    python tools/ncu_target.py --specs fp32:v4:32768:1x1,strict-fma:auto:32768:1x1,fp32:v4:16384:2x1,fp32:v4:16384:1x1:conv

spec = arith:layout:n:PxQ[:conv]   (PxQ > 1x1 = virtual tiles on one GPU: the halo-forwarding
kernel with its peer stores and flags pointed at local buffers)
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from heat2d_b200 import HeatConfig
from heat2d_b200.models.heat import Heat2DSolver

ap = argparse.ArgumentParser()
ap.add_argument("--specs", default="fp32:auto:8192:1x1")
ap.add_argument("--k", type=int, default=8)
ap.add_argument("--launches", type=int, default=1)
a = ap.parse_args()
for spec in a.specs.split(","):
    f = spec.split(":")
    arith, lay, n = f[0], f[1], int(f[2])
    px, py = (int(v) for v in f[3].split("x"))
    conv = len(f) > 4 and f[4] == "conv"
    cfg = HeatConfig(nx=n, ny=n, steps=a.k * a.launches, k=a.k, arith=arith, layout=lay, px=px, py=py,
                     conv=conv, interval=a.k, eps=0.0)
    s = Heat2DSolver(cfg, "cuda:0").initialize()
    r = s.run()
    torch.cuda.synchronize()
    print(f"{spec} k={a.k}: {r.glups:.1f} GLUP/s over {r.launches} launches", flush=True)
    s.close()
    del s
    torch.cuda.empty_cache()
