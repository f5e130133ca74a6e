"""Quick device-timed throughput probe of the single-GPU kernels (not the bench)."""
import argparse
import json
import sys
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from heat2d_b200 import HeatConfig
from heat2d_b200.models.heat import Heat2DSolver


def time_cfg(cfg, reps=3):
    s = Heat2DSolver(cfg, "cuda:0").initialize()
    s.run(max(cfg.k, 1) * 2)  # warm-up
    best = None
    for _ in range(reps):
        r = s.run(cfg.steps)
        g = cfg.nx * cfg.ny * cfg.steps / r.elapsed_s / 1e9
        best = g if best is None else max(best, g)
    s.close()
    del s
    torch.cuda.empty_cache()
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=16384)
    ap.add_argument("--nx", type=int, default=0)
    ap.add_argument("--ny", type=int, default=0)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--ks", default="1,2,4,6,8")
    ap.add_argument("--chunks", default="0")
    ap.add_argument("--arith", default="fp32")
    ap.add_argument("--plain", action="store_true")
    ap.add_argument("--layouts", default="auto", help="comma list of auto,v4,ab2,ab4 (fp32 only)")
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    nx, ny = (a.nx or a.n), (a.ny or a.n)
    out = []
    if a.plain:
        g = time_cfg(HeatConfig(nx=nx, ny=ny, steps=16, backend="plain", k=1, arith=a.arith))
        out.append({"kernel": "plain_k1", "glups": g})
        print(out[-1], flush=True)
    for lay in a.layouts.split(","):
        for k in map(int, a.ks.split(",")):
            for ch in map(int, a.chunks.split(",")):
                g = time_cfg(HeatConfig(nx=nx, ny=ny, steps=a.steps, k=k, chunk_rows=ch, arith=a.arith, layout=lay))
                out.append({"kernel": f"tk_k{k}", "layout": lay, "chunk_rows": ch, "arith": a.arith, "n": [nx, ny],
                            "steps": a.steps, "strict_fast": os.environ.get("H2D_STRICT_FAST", "1"), "glups": g})
                print(out[-1], flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    with open(f"gpurun_out/probe_{a.arith}_{nx}x{ny}{a.tag}.json", "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
