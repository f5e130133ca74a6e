"""How much slower is the halo-forwarding (XTR) kernel than the plain one?  One GPU: the same
grid as one tile and as P x Q virtual tiles (same kernels, peer stores and flags as the
multi-GPU path, pointed at local buffers).

    python tools/xtr_probe.py [n] [steps]
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from heat2d_b200 import HeatConfig
from heat2d_b200.models.heat import Heat2DSolver

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 96
out = []
for arith in os.environ.get("XTR_ARITHS", "fp32,strict-fma").split(","):
    for px, py in ((1, 1), (2, 1), (1, 2)):
        s = Heat2DSolver(HeatConfig(nx=n, ny=n, steps=steps, k=8, px=px, py=py, arith=arith), "cuda:0").initialize()
        s.run(16)
        best = min(s.run(steps).elapsed_s for _ in range(3))
        rec = {"arith": arith, "px": px, "py": py, "n": n, "glups": round(n * n * steps / best / 1e9, 1),
               "fallbacks": [e.strict_fallbacks() for e in s.engines][:2], "launches": s.kernel_launches,
               "chunk_rows": s.engines[0].plan._lib.h2d_plan_chunk_rows_ex(__import__("ctypes").byref(s.fields[0].tile), 8, int(px * py > 1), 0, 1),
               "nofwd": os.environ.get("H2D_DEBUG_NOFWD", "0"), "min_waves": os.environ.get("H2D_MIN_WAVES", "")}
        out.append(rec)
        print(rec, flush=True)
        s.close()
        del s
        torch.cuda.empty_cache()
os.makedirs("gpurun_out", exist_ok=True)
tag = os.environ.get("XTR_TAG", "")
with open(f"gpurun_out/xtr_probe_{n}{tag}.json", "w") as f:
    json.dump(out, f, indent=1)
