"""How far does the fp32 throughput mode drift from the reference arithmetic?

Runs the CPU oracle (bit-identical to the GPU kernels in every mode) in fp32 and in
strict-fma for the same grid and step count and reports max |rel|, max |abs| and how many
cells print differently with the reference's "%6.1f".  Result: profiles/r2/drift_fp32_vs_strict.json
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from heat2d_b200 import _native as N
from heat2d_b200.ops import kernels as K

out = []
for nx, ny, steps in [(2560, 2048, 1000), (16384, 16384, 1000)] if len(sys.argv) < 2 else [tuple(map(int, sys.argv[1:4]))]:
    res = {}
    for arith in ("fp32", "strict-fma"):
        u = K.cpu_initial_grid(nx, ny, "exact")
        t0 = time.time()
        K.cpu_run(u, steps, N.make_phys(0.1, 0.1, "fixed", arith), threads=8)
        res[arith] = u
        print(nx, ny, arith, round(time.time() - t0, 1), "s", flush=True)
    a, b = res["fp32"].astype(np.float64), res["strict-fma"].astype(np.float64)
    nz = b != 0
    rel = np.abs(a[nz] - b[nz]) / np.abs(b[nz])
    # "%6.1f" text: same as rounding to one decimal (ties to even on the exact binary value)
    ta, tb = np.round(a * 10.0), np.round(b * 10.0)
    rec = {"grid": [nx, ny], "steps": steps, "max_rel": float(rel.max()), "mean_rel": float(rel.mean()),
           "max_abs": float(np.abs(a - b).max()), "field_max": float(b.max()),
           "cells_printing_differently_pct": float(100.0 * np.mean(ta != tb)),
           "cells_bit_different_pct": float(100.0 * np.mean(res["fp32"] != res["strict-fma"]))}
    print(rec, flush=True)
    out.append(rec)
os.makedirs(os.path.join(os.path.dirname(__file__), "..", "profiles", "r2"), exist_ok=True)
with open(os.path.join(os.path.dirname(__file__), "..", "profiles", "r2", "drift_fp32_vs_strict.json"), "w") as f:
    json.dump(out, f, indent=1)
