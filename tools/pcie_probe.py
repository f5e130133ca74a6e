"""Host <-> device copy ceiling of this box, one process per GPU (torchrun) or one process.

What the end-to-end number of bench.py can reach at most: every rank copies a pinned buffer
to its GPU, from its GPU, and both at once; rates are per rank and summed over ranks (the
ranks run concurrently: barrier, then time).  Also times the strided (2-D) copies the
pipelined solve issues.  Result: gpurun_out/pcie_probe_n<N>.json (copy into profiles/).

    python tools/pcie_probe.py
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29611 tools/pcie_probe.py
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from heat2d_b200.parallel.dist import init_distributed
from heat2d_b200.utils.numa import bind_to_gpu_numa


def main():
    rank, world, local, device = init_distributed()
    numa = bind_to_gpu_numa(local) if os.environ.get("H2D_NO_NUMA") != "1" else {"bound": False, "skipped": True}
    mib = int(os.environ.get("PCIE_PROBE_MIB", "1024"))
    n = mib * 2**20 // 4
    rows, cols = 8192, n // 8192
    h_in = torch.empty(n, dtype=torch.float32).pin_memory()
    h_out = torch.empty(n, dtype=torch.float32).pin_memory()
    h_in.fill_(1.0)
    h_out.fill_(0.0)                        # first touch after the binding
    d_a = torch.empty(n, dtype=torch.float32, device=device)
    d_b = torch.ones(n, dtype=torch.float32, device=device)
    d_pad = torch.empty((rows, cols + 64), dtype=torch.float32, device=device)
    s1, s2 = torch.cuda.Stream(device), torch.cuda.Stream(device)

    def timed(fn, reps=3):
        best = 1e9
        for _ in range(reps):
            torch.cuda.synchronize(device)
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize(device)
            best = min(best, time.perf_counter() - t0)
        return best

    def h2d():
        with torch.cuda.stream(s1):
            d_a.copy_(h_in, non_blocking=True)

    def d2h():
        with torch.cuda.stream(s2):
            h_out.copy_(d_b, non_blocking=True)

    def both():
        h2d()
        d2h()

    def h2d_2d():
        with torch.cuda.stream(s1):
            d_pad[:, 32:32 + cols].copy_(h_in.view(rows, cols), non_blocking=True)

    def d2h_2d():
        with torch.cuda.stream(s2):
            h_out.view(rows, cols).copy_(d_pad[:, 32:32 + cols], non_blocking=True)

    def both_2d():
        h2d_2d()
        d2h_2d()

    res = {}
    gb = n * 4 / 1e9
    for name, fn, factor in (("h2d", h2d, 1), ("d2h", d2h, 1), ("both", both, 2), ("h2d_2d", h2d_2d, 1),
                             ("d2h_2d", d2h_2d, 1), ("both_2d", both_2d, 2)):
        fn()
        t = timed(fn)
        res[name] = gb * factor / t
    vals = torch.tensor([res[k] for k in sorted(res)], dtype=torch.float64, device=device)
    tot = vals.clone()
    mn = vals.clone()
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
    if rank == 0:
        keys = sorted(res)
        out = {"n_gpus": world, "mib_per_copy": mib, "numa_rank0": numa,
               "gbs_per_rank_min": {k: round(float(v), 2) for k, v in zip(keys, mn.tolist())},
               "gbs_total": {k: round(float(v), 2) for k, v in zip(keys, tot.tolist())},
               "what": "pinned host <-> device copies, all ranks at once; 'both' counts bytes of both directions; "
                       "_2d = row-strided device side (the pipelined solve's copies)"}
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/pcie_probe_n{world}.json", "w") as f:
            json.dump(out, f, indent=1)
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
