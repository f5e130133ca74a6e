import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from heat2d_b200 import HeatConfig
from heat2d_b200.models.heat import Heat2DSolver
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
steps = 1000
s = Heat2DSolver(HeatConfig(nx=n, ny=n, steps=steps, k=8), "cuda:0").initialize()
hin = torch.empty((n, n), dtype=torch.float32).pin_memory(); hin.copy_(s.fields[0].owned(0))
hout = torch.empty((n, n), dtype=torch.float32).pin_memory()
s.run(24)
r = s.run(steps); print("device-only", round(n*n*steps/r.elapsed_s/1e9, 1), flush=True)
for block, skew in [(0, 0), (11448, 0), (15264, 0), (22896, 0), (7632, 12), (15264, 12)]:
    best = 1e9
    for _ in range(2):
        r = s.solve_pipelined(hin, hout, steps, block_rows=block, skew_rows=skew); best = min(best, r.elapsed_s)
    print(f"block={block} skew={skew}: {best*1e3:.1f} ms  {n*n*steps/best/1e9:.1f} GLUP/s  launches {r.launches}", flush=True)
