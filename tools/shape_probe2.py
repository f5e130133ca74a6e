import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from heat2d_b200 import HeatConfig
from heat2d_b200.models.heat import Heat2DSolver
def t(nx, ny, px=1, py=1, steps=64):
    s = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=steps, k=8, px=px, py=py), "cuda:0").initialize()
    s.run(16)
    best = 0
    for _ in range(2):
        r = s.run(steps); best = max(best, nx*ny*steps/r.elapsed_s/1e9)
    # per-rank time of one launch
    import time
    per = []
    for e in s.engines:
        torch.cuda.synchronize(); ev0=torch.cuda.Event(enable_timing=True); ev1=torch.cuda.Event(enable_timing=True)
        ev0.record(); s.epoch += 1; e.launch(s.parity, 8, s.epoch, "none", 0); ev1.record(); ev1.synchronize(); per.append(round(ev0.elapsed_time(ev1)*1e3))
    s.close(); del s; torch.cuda.empty_cache()
    print(f"{nx}x{ny} tiles {px}x{py}: {best:.1f} GLUP/s; per-rank launch us: {per}", flush=True)
t(32768, 32768, 1, 4); t(32768, 32768, 4, 1); t(32768, 32768, 2, 2); t(32768, 32768, 1, 2)
