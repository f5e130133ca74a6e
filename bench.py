#!/usr/bin/env python
"""bench.py — the headline benchmark of heat2d-b200.

Metric (BASELINE.json): grid-point updates per second, GLUP/s = NX*NY*steps / t / 1e9,
device-timed (CUDA events), max over ranks, whole-job aggregate over all N GPUs.

Config (BASELINE.json, config 3): strong scaling of a 32768 x 32768 fp32 grid,
fixed boundary, cx = cy = 0.1, 2-D block decomposition at 1/2/4/8 B200.
``--grid 16384`` gives config 2 (single B200, 16384^2).  A "step" is one time
step of the explicit 5-point update over the whole grid; the fused solver takes
k of them per kernel launch (temporal blocking), every step is really computed.

    python bench.py --gpus 1 --steps 1000 --warmup 24
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29500 bench.py --gpus 8 --steps 1000 --warmup 24
    python bench.py --impl reference --gpus 1 --steps 1000 --warmup 24

The reference arm builds and runs the UNMODIFIED grad1612_cuda_heat.cu staged in
baseline/_ref (see baseline/ref_build.py) for the same grid and step count and
reports the time the reference itself prints.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# Best published throughput of any reference program on its own hardware
# (grad1612 MPI, 2560x2048, 160 ranks, 0.518 s -> 10.1 GLUP/s).
PUBLISHED_BEST_GLUPS = 10.1
PUBLISHED_CUDA_GLUPS = 0.669  # grad1612_cuda_heat, 2560x2048, unnamed GPU


def parse_args():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=24)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "nccl-halo"])
    ap.add_argument("--grid", type=int, default=32768, help="NX = NY (32768: config 3, 16384: config 2)")
    ap.add_argument("--k", type=int, default=8, help="time steps per launch / halo exchange")
    ap.add_argument("--chunk-rows", type=int, default=0)
    ap.add_argument("--arith", default="fp32")
    ap.add_argument("--px", type=int, default=0)
    ap.add_argument("--py", type=int, default=0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--conv", action="store_true", help="config 4: convergence check every 20 steps")
    ap.add_argument("--weak", action="store_true",
                    help="config 5: weak scaling, --grid is the edge of the PER-GPU tile (use --grid 65536)")
    return ap.parse_args()


def reference_arm(a):
    """Run the unmodified reference CUDA program for the same grid / steps."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    out = {"impl": "reference"}
    if rank != 0:
        # the reference has no multi-GPU program: the other ranks have nothing to run
        return None
    try:
        sys.path.insert(0, os.path.join(ROOT, "baseline"))
        import ref_build
        if not ref_build.staged():
            ref_build.stage()
        exe = ref_build.build_cuda(a.grid, a.grid, a.steps)
        warm = ref_build.build_cuda(a.grid, a.grid, max(2, a.warmup - a.warmup % 2))
    except Exception as e:  # noqa: BLE001
        out["unavailable"] = f"cannot stage/build the reference: {str(e).splitlines()[0][:160]}"
        return out
    from heat2d_b200.utils.clocks import ClockSampler
    try:
        ref_build.run_cuda(warm)  # warm-up run (driver, clocks); the program has no warm-up of its own
        with ClockSampler(0, 0.2) as cs:
            secs = ref_build.run_cuda(exe)
    except Exception as e:  # noqa: BLE001
        out["unavailable"] = f"reference run failed: {str(e).splitlines()[0][:160]}"
        return out
    # The reference loop runs k = 0, 2, 4, ... < STEPS with two launches per trip:
    # an odd STEPS performs STEPS + 1 updates (grad1612_cuda_heat.cu:82-85).
    updates = a.steps + (a.steps & 1)
    glups = a.grid * a.grid * updates / secs / 1e9
    out.update({
        "metric": "grid-point updates/sec (GLUP/s, device-timed)", "value": glups, "unit": "GLUP/s",
        "n_gpus": a.gpus, "n_gpus_used": 1, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": secs / updates * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": glups / PUBLISHED_BEST_GLUPS, "dtype": "fp32 storage, fp64 intermediates (as shipped)",
        "data": "synthetic (the reference's own inidat + its zero-field launch-order bug; timing is data-independent)",
        "config": {"model": "grad1612_cuda_heat.cu unmodified, #defines NXPROB/NYPROB/STEPS rewritten, "
                            "nvcc -O3 sm_100a", "grid": [a.grid, a.grid], "global_batch": a.grid * a.grid,
                   "seq_len": a.steps, "parallelism": "1 GPU (the reference has no multi-GPU path)",
                   "l2": "inputs larger than L2"},
        "clocks": cs.summary(), "gpu_launches": updates,
        "note": "time printed by the reference binary (cudaEvent around its launch loop)",
    })
    if world > 1:
        out["note"] += "; ran on GPU 0 only, other ranks idle"
    return out


def main():
    a = parse_args()
    if a.impl == "reference":
        out = reference_arm(a)
        if out is not None:
            print(json.dumps(out), flush=True)
        return 0

    import torch
    import torch.distributed as dist
    from heat2d_b200 import HeatConfig
    from heat2d_b200.parallel.decomp import Decomposition
    from heat2d_b200.parallel.dist import DistributedHeat2DSolver, HaloExchangeSolver, init_distributed
    from heat2d_b200.models.heat import Heat2DSolver
    from heat2d_b200.utils.clocks import ClockSampler

    rank, world, local, device = init_distributed()
    from heat2d_b200.utils.numa import bind_to_gpu_numa
    numa = bind_to_gpu_numa(local)      # pinned host buffers next to the GPU
    if world != a.gpus and world > 1:
        raise SystemExit(f"--gpus {a.gpus} but WORLD_SIZE={world}")
    n = a.grid
    if a.px and a.py:
        px, py = a.px, a.py
    else:
        d = Decomposition.auto(n, n, world)
        px, py = d.px, d.py
    nx, ny = (n * px, n * py) if a.weak else (n, n)
    backend = "nccl-halo" if a.impl == "nccl-halo" else "fused"
    k = 1 if backend == "nccl-halo" else a.k
    cfg = HeatConfig(nx=nx, ny=ny, steps=a.steps, px=px, py=py, k=k, halo=max(k, 8), arith=a.arith,
                     boundary="fixed", init="exact", backend=backend, chunk_rows=a.chunk_rows,
                     conv=a.conv, interval=20, eps=0.0 if a.conv else 0.1, conv_mode="exact")
    if backend == "nccl-halo":
        solver = HaloExchangeSolver(cfg, device)
    elif world > 1:
        solver = DistributedHeat2DSolver(cfg, device)
    else:
        solver = Heat2DSolver(cfg, device)
    solver.initialize()

    def sync_all():
        torch.cuda.synchronize(device)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(device)

    # ---- warm-up (untimed) ----
    if a.warmup > 0:
        solver.run(a.warmup)
    launches0 = solver.kernel_launches
    # ---- timed region: exactly a.steps steps, device-timed, max over ranks ----
    sync_all()
    with ClockSampler(local, 0.1) as cs:
        res = solver.run(a.steps)
        sync_all()
    launches = solver.kernel_launches - launches0
    secs = res.elapsed_s
    glups = nx * ny * a.steps / secs / 1e9

    # ---- end-to-end through the public API: pinned host -> device, K steps, device -> pinned host
    e2e = None
    if not a.no_e2e and backend == "fused":
        f = solver.fields[0]
        t = f.tile
        host_out = torch.empty((t.rows, t.cols), dtype=torch.float32).pin_memory()
        if world > 1:
            host_in = torch.empty(tuple(f.buf[0].shape), dtype=torch.float32).pin_memory()
            host_in.copy_(f.buf[solver.parity])      # any valid field of the right shape
        sync_all()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        if world == 1:
            # one GPU: the public pipelined solve (upload / steps / download overlapped block-wise)
            host_in2 = torch.empty((t.rows, t.cols), dtype=torch.float32).pin_memory()
            host_in2.copy_(f.owned(solver.parity))
            torch.cuda.synchronize(device)
            t0 = time.perf_counter()
            ev0.record()
            solver.solve_pipelined(host_in2, host_out, a.steps)
            ev1.record()
            ev1.synchronize()
            host_in = host_in2
            e2e_how = "Heat2DSolver.solve_pipelined: upload, steps and download overlapped block by block"
        elif os.environ.get("H2D_BENCH_E2E", "legacy") == "pipelined" and a.impl == "ours" and not a.conv \
                and not a.weak and ny % (4 * world) == 0 and ny // world >= 64:
            # several GPUs (opt-in, H2D_BENCH_E2E=pipelined: validated on one GPU with virtual tiles only):
            # column strips, every rank pipelines upload / windows / download in lockstep
            scfg = cfg.replace(px=1, py=world)
            strip = DistributedHeat2DSolver(scfg, device).initialize()
            hin, col0, host_out = strip.host_strip()
            sf = strip.fields[0]
            st = sf.tile
            hin.copy_(sf.buf[0][st.halo_r:st.halo_r + st.rows, st.halo_c - col0:st.halo_c - col0 + hin.shape[1]])
            strip.barrier()
            t0 = time.perf_counter()
            ev0.record()
            strip.solve_pipelined(hin, col0, host_out, a.steps)
            ev1.record()
            ev1.synchronize()
            host_in = hin
            e2e_how = ("DistributedHeat2DSolver.solve_pipelined on a 1x%d grid: per rank upload, window launches with "
                       "W/E halo exchange and download overlapped block by block, ranks in lockstep" % world)
            strip.close()
        else:
            ev0.record()
            f.buf[solver.parity].copy_(host_in, non_blocking=True)       # H2D of this rank's inputs
            solver.barrier()                                             # peers' ghosts must be in place
            r2 = solver.run(a.steps)
            host_out.copy_(f.owned(solver.parity), non_blocking=True)    # D2H of the result
            ev1.record()
            ev1.synchronize()
            e2e_how = "per rank: upload of the padded tile, barrier, run(), download of the owned cells (not overlapped)"
        wall = time.perf_counter() - t0
        e2e_s = max(ev0.elapsed_time(ev1) / 1e3, 0.0)
        tt = torch.tensor([e2e_s, wall], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_s = float(tt[0].item())
        e2e = {"value": nx * ny * a.steps / e2e_s / 1e9, "unit": "GLUP/s",
               "h2d_bytes_per_step": host_in.numel() * 4 * world / a.steps,
               "d2h_bytes_per_step": host_out.numel() * 4 * world / a.steps,
               "seconds": e2e_s, "wall_seconds": float(tt[1].item()),
               "what": "pinned host -> device copy of the whole initial field, "
                       f"{a.steps} steps, device -> pinned host copy of the whole final field; bytes are "
                       "the job totals divided by the step count", "how": e2e_how}

    if rank == 0:
        out = {
            "metric": "grid-point updates/sec (GLUP/s, device-timed, max over ranks)",
            "value": glups, "unit": "GLUP/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": secs / a.steps * 1e3, "higher_is_better": True,
            "scaling": "weak" if a.weak else "strong",
            "vs_baseline": glups / PUBLISHED_BEST_GLUPS, "dtype": "fp32" if a.arith == "fp32" else a.arith,
            "data": "synthetic (reference initial condition x(NX-x-1)y(NY-y-1), exact int64 product)",
            "config": {"model": "heat2d 5-point explicit stencil, fixed boundary, cx=cy=0.1",
                       "grid": [nx, ny], "global_batch": nx * ny, "seq_len": a.steps,
                       "parallelism": f"{px}x{py} tiles (rows x cols), one per GPU",
                       "k": k, "backend": a.impl, "conv": bool(a.conv),
                       "l2": f"inputs larger than L2 ({nx * ny * 4 / world / 2**20:.0f} MiB per GPU per buffer vs 126 MiB)"},
            "clocks": cs.summary(), "gpu_launches": launches, "impl": a.impl, "numa": numa,
            "halo_wait_us_rank0": (solver.halo_wait_stats()[0] if hasattr(solver, "halo_wait_stats") else None),
            "baseline_ref": "10.1 GLUP/s = best published reference number (grad1612 MPI, 2560x2048, 160 ranks, "
                            "Report.pdf p21); published CUDA number is 0.669 GLUP/s",
            "roofline": _roofline(glups, nx, ny, px, py, k, a.chunk_rows),
        }
        if e2e is not None:
            out["e2e"] = e2e
        print(json.dumps(out), flush=True)
    close = getattr(solver, "close", None)
    if close:
        close()
    if world > 1 and dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    return 0


def _roofline(glups, nx, ny, px, py, k, chunk_rows):
    """Achieved fraction of the three-roof bound (fp32 pipe / HBM / NVLink), whole job."""
    out = {"hbm_gbs_measured": _peaks().get("hbm_gbs"),
           "k1_hbm_bound_glups_per_gpu": (_peaks().get("hbm_gbs") or 0) / 8.0}
    try:
        import ctypes
        from heat2d_b200 import _native as N
        from heat2d_b200.utils.perfmodel import roofline_fraction
        rows, cols = -(-nx // px), -(-ny // py)
        if chunk_rows <= 0:
            t = N.Tile(nx, ny, 0, 0, rows, cols, max(k, 8), 32, cols + 64)
            chunk_rows = int(N.cuda().h2d_plan_chunk_rows(ctypes.byref(t), k, int(px * py > 1)))
        out.update(roofline_fraction(glups, nx, ny, px, py, k, chunk_rows))
        out["chunk_rows"] = chunk_rows
    except Exception as e:  # noqa: BLE001 - the model is a report, never a reason to lose the bench line
        out["model_error"] = str(e)[:120]
    return out


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except Exception:  # noqa: BLE001
        return {}


if __name__ == "__main__":
    sys.exit(main())
