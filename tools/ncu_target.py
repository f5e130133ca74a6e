"""Small program for ncu captures: a few launches of one kernel variant."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from heat2d_b200 import HeatConfig
from heat2d_b200.models.heat import Heat2DSolver

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=8192)
ap.add_argument("--k", type=int, default=8)
ap.add_argument("--launches", type=int, default=4)
ap.add_argument("--backend", default="fused")
ap.add_argument("--arith", default="fp32")
ap.add_argument("--chunk-rows", type=int, default=0)
a = ap.parse_args()
cfg = HeatConfig(nx=a.n, ny=a.n, steps=a.k * a.launches, k=a.k, backend=a.backend, arith=a.arith,
                 chunk_rows=a.chunk_rows)
s = Heat2DSolver(cfg, "cuda:0").initialize()
r = s.run()
torch.cuda.synchronize()
print(f"{a.backend} k={a.k} n={a.n}: {r.glups:.1f} GLUP/s over {r.launches} launches")
