"""Sweep the block / skew parameters of the pipelined upload-solve-download on one GPU.

    python tools/pipelined_probe.py [n] [steps] [arith]
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from heat2d_b200 import HeatConfig
from heat2d_b200.models.heat import Heat2DSolver

n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
arith = sys.argv[3] if len(sys.argv) > 3 else "fp32"
s = Heat2DSolver(HeatConfig(nx=n, ny=n, steps=steps, k=8, arith=arith), "cuda:0").initialize()
hin = torch.empty((n, n), dtype=torch.float32).pin_memory()
hin.copy_(s.fields[0].owned(0))
hout = torch.empty((n, n), dtype=torch.float32).pin_memory()
s.run(24)
r = s.run(steps)
out = {"n": n, "steps": steps, "arith": arith, "device_only_glups": round(n * n * steps / r.elapsed_s / 1e9, 1), "sweep": []}
print("device-only", out["device_only_glups"], flush=True)
wave = s.engines[0].plan.wave_rows
for block, skew in [(0, 0), (wave // 2, 0), (wave, 0), (2 * wave, 0), (3 * wave, 0), (6 * wave, 0), (wave, 9), (2 * wave, 9)]:
    best = 1e9
    for _ in range(3):
        r = s.solve_pipelined(hin, hout, steps, block_rows=block, skew_rows=skew)
        best = min(best, r.elapsed_s)
    out["sweep"].append({"block_rows": block, "skew_rows": skew, "ms": round(best * 1e3, 2),
                         "glups": round(n * n * steps / best / 1e9, 1), "launches": r.launches})
    print(out["sweep"][-1], flush=True)
os.makedirs("gpurun_out", exist_ok=True)
with open(f"gpurun_out/pipelined_probe_{n}_{steps}_{arith}.json", "w") as f:
    json.dump(out, f, indent=1)
