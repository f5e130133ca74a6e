"""The six grids of the reference's report (Table 10), 1000 steps, on this B200:
unmodified grad1612_cuda_heat.cu (rebuilt for sm_100a) vs this framework."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "baseline"))
import torch

import ref_build
from heat2d_b200 import HeatConfig
from heat2d_b200.models.heat import Heat2DSolver
from heat2d_b200.ops import kernels as K

PUBLISHED = {(80, 64): 9.15e-3, (160, 128): 2.19e-2, (320, 256): 7.14e-2, (640, 512): 0.327,
             (1280, 1024): 1.86, (2560, 2048): 7.84}
rows = []
for (nx, ny), pub in PUBLISHED.items():
    rec = {"grid": [nx, ny], "steps": 1000, "published_s": pub}
    try:
        exe = ref_build.build_cuda(nx, ny, 1000)
        ref_build.run_cuda(exe)
        rec["reference_b200_s"] = min(ref_build.run_cuda(exe) for _ in range(3))
    except Exception as e:  # noqa: BLE001
        rec["reference_b200_s"] = None
        rec["reference_error"] = str(e)[:100]
    best = None
    for backend in ("fused", "resident"):
        if backend == "resident" and not K.resident_fits(nx, ny):
            continue
        s = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=1000, k=8, backend=backend), "cuda:0").initialize()
        s.run(64)
        t = min(s.run(1000).elapsed_s for _ in range(3))
        rec[f"ours_{backend}_s"] = t
        best = t if best is None else min(best, t)
        s.close()
    rec["ours_best_s"] = best
    rec["glups_ours"] = nx * ny * 1000 / best / 1e9
    if rec["reference_b200_s"]:
        rec["glups_reference_b200"] = nx * ny * 1000 / rec["reference_b200_s"] / 1e9
        rec["speedup_vs_reference_b200"] = rec["reference_b200_s"] / best
    rec["speedup_vs_published"] = pub / best
    rows.append(rec)
    print(rec, flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/published_sizes.json", "w"), indent=1)
