"""Cluster (team) launch vs the independent-strip launch of the temporal-blocked kernel, one GPU:
same field, same k, bit-compared, device-timed.

    python tools/team_probe.py [--n 32768] [--launches 50] [--ariths fp32,strict-fma] [--clusters 1,2,4]

Writes gpurun_out/team_probe_<n>.json.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from heat2d_b200 import HeatConfig
from heat2d_b200.models.heat import Heat2DSolver


def timed(fn, launches, reps=3):
    best = None
    for _ in range(reps):
        t0 = torch.cuda.Event(enable_timing=True)
        t1 = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        t0.record()
        par = 0
        for _ in range(launches):
            par = fn(par)
        t1.record()
        t1.synchronize()
        ms = t0.elapsed_time(t1)
        best = ms if best is None else min(best, ms)
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=32768)
    ap.add_argument("--launches", type=int, default=50)
    ap.add_argument("--ariths", default="fp32,strict-fma")
    ap.add_argument("--clusters", default="1,2,4")
    ap.add_argument("--ks", default="8")
    ap.add_argument("--chunks", default="0")
    ap.add_argument("--checks", type=int, default=4)
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    n = a.n
    out = []
    for arith in a.ariths.split(","):
        for k in map(int, a.ks.split(",")):
            s = Heat2DSolver(HeatConfig(nx=n, ny=n, steps=k, k=k, arith=arith, layout="v4"), "cuda:0").initialize()
            e = s.engines[0]
            f = s.fields[0]
            init = f.buf[0].clone()
            # reference result of one launch
            e.plan.launch(0, k, 1, "none", 0, 0)
            torch.cuda.synchronize()
            want = f.owned(1).clone()

            def base(par):
                e.plan.launch(par, k, 1, "none", 0, 0)
                return par ^ 1
            base(0), base(1)
            ms = timed(base, a.launches)
            rec = {"arith": arith, "k": k, "n": n, "launch": "independent strips",
                   "glups": round(n * n * k * a.launches / ms / 1e6, 1), "ms_per_launch": round(ms / a.launches, 4)}
            out.append(rec)
            print(rec, flush=True)
            for cl in map(int, a.clusters.split(",")):
                for ch in map(int, a.chunks.split(",")):
                    same = True
                    for _ in range(a.checks):
                        f.buf[0].copy_(init)
                        f.buf[1].zero_()
                        e.plan.launch_team(0, k, cl, ch)
                        torch.cuda.synchronize()
                        same = same and bool(torch.equal(f.owned(1), want))

                    def team(par, cl=cl, ch=ch):
                        return e.plan.launch_team(par, k, cl, ch)
                    team(0), team(1)
                    ms = timed(team, a.launches)
                    rec = {"arith": arith, "k": k, "n": n, "launch": f"teams, cluster of {cl}", "chunk_rows": ch,
                           "bit_identical": same, "glups": round(n * n * k * a.launches / ms / 1e6, 1),
                           "ms_per_launch": round(ms / a.launches, 4)}
                    out.append(rec)
                    print(rec, flush=True)
            s.close()
            del s, e, f, init, want
            torch.cuda.empty_cache()
    os.makedirs("gpurun_out", exist_ok=True)
    with open(f"gpurun_out/team_probe_{n}{a.tag}.json", "w") as fo:
        json.dump(out, fo, indent=1)


if __name__ == "__main__":
    main()
