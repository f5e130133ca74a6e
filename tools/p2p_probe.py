"""Measure what the roofline needs and MEASURED_PEAKS.json does not have:
peer-to-peer store bandwidth over NVLink (kernel stores into a mapped peer arena,
the same mechanism the fused halo exchange uses) and the small-message latency of
a flag round trip.  Run with torchrun on >= 2 GPUs."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from heat2d_b200 import _native as N
from heat2d_b200.models.heat import tensor_from_ptr
from heat2d_b200.parallel.dist import init_distributed


def main():
    rank, world, local, device = init_distributed("nccl")
    lib = N.cuda()
    nbytes = 1 << 30
    h = lib.h2d_arena_create(nbytes)
    ptr = int(lib.h2d_arena_ptr(h))
    import ctypes as C
    buf = C.create_string_buffer(64)
    N.check(lib, lib.h2d_arena_export(h, buf))
    handles = [None] * world
    dist.all_gather_object(handles, buf.raw)
    peer = (rank + 1) % world
    pp = int(lib.h2d_arena_import(handles[peer]))
    mine = tensor_from_ptr(ptr, (nbytes // 4,), torch.float32, device)
    theirs = tensor_from_ptr(pp, (nbytes // 4,), torch.float32, device)
    src = torch.ones(nbytes // 4, dtype=torch.float32, device=device)
    out = {}
    for name, n in (("1GiB", nbytes // 4), ("64MiB", (64 << 20) // 4), ("1MiB", (1 << 20) // 4), ("64KiB", (64 << 10) // 4)):
        for _ in range(3):
            theirs[:n].copy_(src[:n])
        torch.cuda.synchronize()
        dist.barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 10 if n > (1 << 22) else 100
        ev0.record()
        for _ in range(reps):
            theirs[:n].copy_(src[:n])      # elementwise kernel: plain st.global on peer addresses
        ev1.record()
        ev1.synchronize()
        ms = ev0.elapsed_time(ev1) / reps
        out[f"store_{name}_GBps"] = n * 4 / ms / 1e6
        out[f"store_{name}_us"] = ms * 1e3
    # peer read
    dst = torch.empty(nbytes // 4, dtype=torch.float32, device=device)
    for _ in range(3):
        dst.copy_(theirs)
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(10):
        dst.copy_(theirs)
    ev1.record()
    ev1.synchronize()
    out["load_1GiB_GBps"] = nbytes / (ev0.elapsed_time(ev1) / 10) / 1e6
    res = [None] * world
    dist.all_gather_object(res, out)
    if rank == 0:
        agg = {k: min(r[k] for r in res) if k.endswith("GBps") else max(r[k] for r in res) for k in out}
        agg["world"] = world
        agg["what"] = "kernel stores/loads on CUDA-IPC mapped peer memory, every rank -> (rank+1) simultaneously; min over ranks"
        print(json.dumps(agg))
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/p2p_probe_w{world}.json", "w") as f:
            json.dump(agg, f, indent=1)
    dist.barrier()
    lib.h2d_arena_unmap(pp)
    dist.barrier()
    lib.h2d_arena_destroy(h)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
