"""Launch-per-sweep vs dataflow launch on one GPU (device-timed, CUDA events).

    python tools/dataflow_probe.py [--grid 32768] [--steps 1000] [--chunks 0,1024,2048]
Writes gpurun_out/dataflow_probe.json."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from heat2d_b200.config import HeatConfig  # noqa: E402
from heat2d_b200.models.heat import Heat2DSolver  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", nargs="+", default=["32768"], help="N or NXxNY")
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--k", type=int, default=8)
    ap.add_argument("--chunks", default="0,1024,2048")
    ap.add_argument("--reps", type=int, default=2)
    a = ap.parse_args()
    out = []
    for spec in a.grid:
        nx, ny = (int(v) for v in spec.split("x")) if "x" in spec else (int(spec), int(spec))
        for dataflow in (False, True):
            for chunk in ([0] if not dataflow else [int(c) for c in a.chunks.split(",")]):
                cfg = HeatConfig(nx=nx, ny=ny, steps=a.steps, k=a.k, chunk_rows=chunk, dataflow=dataflow)
                s = Heat2DSolver(cfg, "cuda:0").initialize()
                s.run(24 * a.k // 8)   # warm-up
                best = None
                for _ in range(a.reps):
                    r = s.run()
                    best = r if best is None or r.elapsed_s < best.elapsed_s else best
                rec = {"grid": [nx, ny], "steps": a.steps, "k": a.k, "dataflow": dataflow, "chunk_rows": chunk,
                       "seconds": best.elapsed_s, "glups": best.glups, "launches": best.launches}
                print(json.dumps(rec), flush=True)
                out.append(rec)
                s.close()
                del s
                torch.cuda.empty_cache()
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/dataflow_probe.json", "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
