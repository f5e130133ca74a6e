"""Small program for ncu captures: a few launches of one or several kernel variants.

    python tools/ncu_target.py --n 16384 --variants fp32:v4,fp32:ab4,strict-fma:auto --launches 2
"""
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
ap.add_argument("--variants", default="fp32:auto", help="comma list of arith:layout")
ap.add_argument("--chunk-rows", type=int, default=0)
ap.add_argument("--px", type=int, default=1)
ap.add_argument("--py", type=int, default=1)
ap.add_argument("--conv", action="store_true")
a = ap.parse_args()
for v in a.variants.split(","):
    arith, lay = v.split(":")
    cfg = HeatConfig(nx=a.n, ny=a.n, steps=a.k * a.launches, k=a.k, backend=a.backend, arith=arith, layout=lay,
                     chunk_rows=a.chunk_rows, px=a.px, py=a.py, conv=a.conv, interval=a.k, eps=0.0)
    s = Heat2DSolver(cfg, "cuda:0").initialize()
    r = s.run()
    torch.cuda.synchronize()
    print(f"{a.backend} {v} k={a.k} n={a.n} {a.px}x{a.py}: {r.glups:.1f} GLUP/s over {r.launches} launches", flush=True)
    s.close()
    del s
    torch.cuda.empty_cache()
