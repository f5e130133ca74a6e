import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from heat2d_b200 import HeatConfig
from heat2d_b200.models.heat import Heat2DSolver
def t(nx, ny, px=1, py=1, steps=64):
    s = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=steps, k=8, px=px, py=py), "cuda:0").initialize()
    s.run(16)
    best = 0
    for _ in range(3):
        r = s.run(steps); best = max(best, nx*ny*steps/r.elapsed_s/1e9)
    s.close(); del s; torch.cuda.empty_cache()
    print(f"{nx}x{ny} tiles {px}x{py}: {best:.1f} GLUP/s", flush=True)
t(32768, 8192); t(8192, 32768)
t(32768, 16384, 1, 2); t(32768, 16384, 2, 1); t(32768, 16384, 1, 1)
t(16384, 16384, 2, 2)
