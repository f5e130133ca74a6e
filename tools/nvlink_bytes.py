"""Measured NVLink traffic of the fused halo exchange: `nvidia-smi nvlink -gt d` (cumulative
Tx / Rx KiB per link) before and after a timed run on N GPUs, against the algorithmic
2 * k * NY * 4 bytes per interior neighbour pair and launch.  Stands in for ncu's NVLink
counters, which cannot be collected on a multi-rank run (B200_PROFILING.md).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29650 tools/nvlink_bytes.py --grid 32768 --steps 400 --px 2 --py 1
"""
import argparse
import json
import os
import re
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from heat2d_b200 import HeatConfig
from heat2d_b200.parallel.dist import DistributedHeat2DSolver, init_distributed


def nvlink_kib(gpu: int):
    """(tx_kib, rx_kib) summed over the links of one GPU."""
    out = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", str(gpu)], capture_output=True, text=True).stdout
    tx = sum(int(v) for v in re.findall(r"Data Tx:\s*(\d+)\s*KiB", out))
    rx = sum(int(v) for v in re.findall(r"Data Rx:\s*(\d+)\s*KiB", out))
    return tx, rx, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=32768)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--k", type=int, default=8)
    ap.add_argument("--px", type=int, default=0)
    ap.add_argument("--py", type=int, default=0)
    ap.add_argument("--arith", default="fp32")
    a = ap.parse_args()
    rank, world, local, device = init_distributed()
    px, py = (a.px, a.py) if a.px and a.py else (world, 1)
    n = a.grid
    s = DistributedHeat2DSolver(HeatConfig(nx=n, ny=n, steps=a.steps, px=px, py=py, k=a.k, arith=a.arith), device).initialize()
    s.run(a.k * 3)
    torch.cuda.synchronize()
    dist.barrier()
    t0, r0, raw0 = nvlink_kib(local)
    dist.barrier()
    res = s.run(a.steps)
    torch.cuda.synchronize()
    dist.barrier()
    t1, r1, raw1 = nvlink_kib(local)
    spec = s.decomp.tile(rank)
    launches = -(-a.steps // a.k)
    nb = spec.neighbours
    expect = 0
    for d in "NS":
        if nb[d] is not None:
            expect += a.k * spec.cols * 4 * launches
    for d in "WE":
        if nb[d] is not None:
            expect += a.k * spec.rows * 4 * launches
    for d in ("NW", "NE", "SW", "SE"):
        if nb[d] is not None:
            expect += a.k * a.k * 4 * launches
    rec = {"rank": rank, "tx_bytes": (t1 - t0) * 1024, "rx_bytes": (r1 - r0) * 1024, "expected_tx_bytes": expect,
           "launches": launches, "counter_lines": len(re.findall(r"Data Tx", raw1))}
    recs = [None] * world
    dist.all_gather_object(recs, rec)
    if rank == 0:
        out = {"n_gpus": world, "grid": n, "steps": a.steps, "k": a.k, "layout": f"{px}x{py}", "arith": a.arith,
               "glups": n * n * a.steps / res.elapsed_s / 1e9, "per_rank": recs,
               "what": "nvidia-smi nvlink -gt d deltas around the timed run; expected = k-wide bands of owned cells per "
                       "neighbour per launch (the flags, 4 bytes per neighbour per launch, and NVLink protocol overhead "
                       "come on top)"}
        if not recs[0]["counter_lines"]:
            out["note"] = "nvidia-smi printed no NVLink data counters on this box: " + raw1[:200].replace("\n", " | ")
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/nvlink_bytes_n{world}_{px}x{py}.json", "w") as f:
            json.dump(out, f, indent=1)
        print(json.dumps(out)[:1500], flush=True)
    s.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
