"""Several multi-GPU configurations in ONE torchrun session (process start-up and NCCL
bootstrap are paid once): process-grid layouts x H2D_MIN_WAVES x arithmetic, device-timed,
max over ranks.  Every configuration is first checked against the CPU oracle on a small grid.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29700 tools/scaling_sweep.py --grid 32768 --steps 1000 \
        --layouts 8x1,4x2,2x4,1x8 --min-waves 1,2,4 --ariths fp32,strict-fma

Writes gpurun_out/scaling_sweep_n<N>.jsonl (one JSON object per line).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from heat2d_b200 import HeatConfig, _native as N
from heat2d_b200.ops import kernels as K
from heat2d_b200.parallel.dist import DistributedHeat2DSolver, HaloExchangeSolver, init_distributed


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=32768)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=24)
    ap.add_argument("--k", type=int, default=8)
    ap.add_argument("--ks", default="", help="comma list of sweep depths (overrides --k)")
    ap.add_argument("--layouts", default="")
    ap.add_argument("--min-waves", default="4")
    ap.add_argument("--ariths", default="fp32")
    ap.add_argument("--backends", default="fused")
    ap.add_argument("--conv", action="store_true")
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--tag", default="")
    ap.add_argument("--bulk", default="1", help="comma list of 0/1: row bands through the bulk-copy engine")
    a = ap.parse_args()
    rank, world, local, device = init_distributed()
    layouts = [tuple(int(v) for v in l.split("x")) for l in a.layouts.split(",") if l] or [(world, 1)]
    out_path = f"gpurun_out/scaling_sweep_n{world}{a.tag}.jsonl"
    if rank == 0:
        os.makedirs("gpurun_out", exist_ok=True)
    n = a.grid
    ks = [int(v) for v in a.ks.split(",") if v] or [a.k]
    for backend in [(b, kk) for b in a.backends.split(",") for kk in ks]:
        backend, a.k = backend
        for arith in a.ariths.split(","):
            for px, py in layouts:
                if px * py != world:
                    continue
                for mw, bulk in [(m, b) for m in a.min_waves.split(",") for b in a.bulk.split(",")]:
                    os.environ["H2D_MIN_WAVES"] = mw
                    os.environ["H2D_FWD_BULK"] = bulk
                    cls = HaloExchangeSolver if backend == "nccl-halo" else DistributedHeat2DSolver
                    rec = {"n_gpus": world, "grid": n, "steps": a.steps, "k": a.k, "backend": backend, "arith": arith,
                           "px": px, "py": py, "min_waves": int(mw), "bulk": int(bulk), "conv": a.conv}
                    try:
                        # correctness on a small case with the same process grid
                        vcfg = HeatConfig(nx=1536, ny=2048, steps=2 * a.k + 3, px=px, py=py, k=a.k, halo=max(a.k, 8),
                                          arith=arith, backend=backend)
                        u0 = (np.random.default_rng(5).random((1536, 2048), dtype=np.float32) * 100).astype(np.float32)
                        s = cls(vcfg, device)
                        s.load(torch.from_numpy(u0))
                        s.run()
                        got = s.gather()
                        ok = True
                        if rank == 0:
                            want = u0.copy()
                            K.cpu_run(want, vcfg.steps, N.make_phys(0.1, 0.1, "fixed", arith), threads=8)
                            ok = bool(np.array_equal(got.cpu().numpy(), want))
                        s.close()
                        cfg = HeatConfig(nx=n, ny=n, steps=a.steps, px=px, py=py, k=a.k, halo=max(a.k, 8), arith=arith,
                                         backend=backend, conv=a.conv, interval=20, eps=0.0 if a.conv else 0.1)
                        s = cls(cfg, device).initialize()
                        s.run(a.warmup)
                        best = None
                        for _ in range(a.reps):
                            r = s.run(a.steps)
                            best = r.elapsed_s if best is None else min(best, r.elapsed_s)
                        wait = s.halo_wait_stats() if hasattr(s, "halo_wait_stats") else (None, None)
                        rec.update({"verified": ok, "glups": n * n * a.steps / best / 1e9, "ms": best * 1e3,
                                    "halo_wait_us_rank0": wait[0], "halo_waits_rank0": wait[1],
                                    "chunk_rows": (s.engines[0].plan._lib.h2d_plan_chunk_rows_ex(
                                        __import__("ctypes").byref(s.fields[0].tile), a.k, 1, N.ARITH[arith], 0)
                                        if backend == "fused" else None)})
                        s.close()
                    except Exception as e:  # noqa: BLE001 - keep sweeping
                        rec["error"] = str(e)[:200]
                    if rank == 0:
                        print(json.dumps(rec), flush=True)
                        with open(out_path, "a") as f:
                            f.write(json.dumps(rec) + "\n")
                    dist.barrier()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
