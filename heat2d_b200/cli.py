"""Command line: the reference's compile-time macros as runtime flags.

    python -m heat2d_b200 --nx 2560 --ny 2048 --steps 1000 -k 8
    torchrun --nproc-per-node 8 -m heat2d_b200 --nx 32768 --ny 32768 --px 2 --py 4 --steps 1000
    python -m heat2d_b200 --preset grad1612_mpi --backend cpu --out-dir out

Stdout keeps the reference's lines (problem size, per-rank tile, iteration count,
``Elapsed time: %e sec``); a JSON record goes to ``<out-dir>/result.json``.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

from .config import BACKENDS, CONV_MODES, HeatConfig, ConfigError


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(prog="heat2d_b200", description=__doc__,
                                 formatter_class=argparse.RawDescriptionHelpFormatter)
    d = HeatConfig()
    ap.add_argument("--preset", choices=["mpi_heat2Dn", "grad1612_mpi", "grad1612_hybrid", "grad1612_cuda"],
                    help="start from one reference program's #define block")
    ap.add_argument("--nx", type=int, help=f"NXPROB: rows (default {d.nx})")
    ap.add_argument("--ny", type=int, help=f"NYPROB: columns (default {d.ny})")
    ap.add_argument("--steps", type=int, help=f"STEPS (default {d.steps})")
    ap.add_argument("--px", type=int, help="GRIDX: tiles along x (rows)")
    ap.add_argument("--py", type=int, help="GRIDY: tiles along y (columns)")
    ap.add_argument("--conv", choices=["on", "off"], help="CONVERGENCE")
    ap.add_argument("--interval", type=int, help="INTERVAL (default 20)")
    ap.add_argument("--eps", type=float, help="SENSITIVITY (default 0.1)")
    ap.add_argument("--norm", choices=["l2sq", "linf"], help="l2sq = reference; linf = max |du|")
    ap.add_argument("--conv-mode", choices=CONV_MODES)
    ap.add_argument("--ref-conv-bug", action="store_true",
                    help="replay the shipped predicate `i %% INTERVAL` (grad1612_mpi_heat.c:263): "
                         "checks every step iff (nx/px) %% interval == 0, else never")
    ap.add_argument("--cx", type=float)
    ap.add_argument("--cy", type=float)
    ap.add_argument("--cx-float", action="store_true", help="coefficients stored as float (mpi_heat2Dn.c:41-44)")
    ap.add_argument("--boundary", choices=["fixed", "ghost-ring"])
    ap.add_argument("--arith", choices=["fp32", "strict", "strict-fma"])
    ap.add_argument("--init", choices=["ref-int32", "ref-uint32", "exact", "zero"])
    ap.add_argument("--dump", help="comma list of none|prtdat|rowmajor|binary")
    ap.add_argument("--out-dir")
    ap.add_argument("-k", type=int, help="time steps per launch / halo exchange (1..8); default: 8, or — without "
                                         "a preset — what is fastest for the arithmetic and tile size (config.best_k)")
    ap.add_argument("--backend", choices=BACKENDS)
    ap.add_argument("--chunk-rows", type=int)
    ap.add_argument("--layout", choices=["auto", "v4", "ab2", "ab4", "v8"],
                    help="lane layout of the kernels (same bits): v4 | ab2 | ab4 for fp32 (v4 is the measured best on "
                         "B200), v4 | v8 for strict arithmetic with -k <= 4")
    ap.add_argument("--team", choices=["on", "off"],
                    help="on: k >= 5 sweeps of single-GPU runs as cluster launches that share edge columns between "
                         "strips instead of recomputing them (same bits; pays off in strict arithmetic at >= 32768^2)")
    ap.add_argument("--dataflow", action="store_true",
                    help="single tile, no convergence: all sweeps in one dataflow launch")
    ap.add_argument("--threads", type=int, help="NUMTHREADS (cpu backend)")
    ap.add_argument("--no-uneven", action="store_true", help="insist on exact divisibility (grad1612_mpi_heat.c:60)")
    ap.add_argument("--restart", help="raw fp32 nx*ny file (final_binary.dat) to start from")
    ap.add_argument("--start-step", type=int, help="step index the restart file corresponds to")
    ap.add_argument("--debug", action="store_true", help="DEBUG: neighbours per rank, device properties")
    return ap


def config_from_args(a) -> HeatConfig:
    if a.preset:
        from .models.presets import preset
        cfg = preset(a.preset)
    else:
        cfg = HeatConfig()
    over = {}
    for name in ("nx", "ny", "steps", "px", "py", "interval", "eps", "norm", "cx", "cy", "boundary", "arith",
                 "init", "dump", "out_dir", "k", "backend", "chunk_rows", "threads", "restart", "start_step",
                 "conv_mode", "layout", "team"):
        v = getattr(a, name, None)
        if v is not None:
            over[name] = v
    if a.conv is not None:
        over["conv"] = a.conv == "on"
    if a.cx_float:
        over["cx_float"] = True
    if a.no_uneven:
        over["uneven"] = False
    if a.debug:
        over["debug"] = True
    if getattr(a, "dataflow", False):
        over["dataflow"] = True
    if a.ref_conv_bug:
        over["ref_conv_bug"] = True
    cfg = cfg.replace(**over)
    if a.k is None and not a.preset:
        from .config import best_k
        cfg = cfg.replace(k=best_k(cfg.arith, (cfg.nx * cfg.ny) // max(cfg.px * cfg.py, 1)))
    if cfg.halo < cfg.k:
        cfg = cfg.replace(halo=cfg.k)
    return cfg


def _banner(cfg: HeatConfig, world: int, style: str):
    if style == "mpi_heat2Dn":
        print(f"Starting mpi_heat2D with {world} worker tasks.")
        print(f"Grid size: X= {cfg.nx}  Y= {cfg.ny}  Time steps= {cfg.steps}")
    elif style == "grad1612_cuda":
        print(f"Problem size: {cfg.nx}x{cfg.ny}\nAmount of iterations: {cfg.steps}")
    else:
        threads = f" and {cfg.threads} threads" if style == "grad1612_hybrid" else ""   # grad1612_hybrid_heat.c:70
        print(f"Starting with {world} processes{threads}\nProblem size:{cfg.nx}x{cfg.ny}\n"
              f"Each process will take: {cfg.nx // cfg.px}x{cfg.ny // cfg.py}\nAmount of iterations: {cfg.steps}")
        if cfg.conv:
            print(f"Check for convergence every {cfg.interval} iterations")


def main(argv=None) -> int:
    a = build_parser().parse_args(argv)
    try:
        cfg = config_from_args(a)
        dumps = [d for d in cfg.dump.split(",") if d and d != "none"]
        for dname in dumps:
            if dname not in ("prtdat", "rowmajor", "binary"):
                raise ConfigError(f"unknown dump format {dname!r}")
        cfg = cfg.replace(dump="none")  # dumps are handled here, formats validated above
        if a.preset == "mpi_heat2Dn" and not 3 <= cfg.px * cfg.py <= 8 and cfg.backend != "cpu":
            # MINWORKER / MAXWORKER of the given program (mpi_heat2Dn.c:32-33,72-78); the master
            # of the reference does no compute, so "workers" are this framework's ranks
            raise ConfigError("ERROR: the number of tasks must be between 4 and 9.")
        if cfg.ref_conv_bug and cfg.conv:
            # the stale loop variable equals the tile height after the interior sweep
            fires = (cfg.nx // cfg.px) % cfg.interval == 0
            cfg = cfg.replace(interval=1) if fires else cfg.replace(conv=False)
    except (ConfigError, KeyError, ValueError) as e:
        print(str(e), file=sys.stderr)
        print("Quitting...", file=sys.stderr)
        return 1

    from ._lazy import LazyModule
    torch = LazyModule("torch")   # the cpu backend never touches it: start-up in ~0.3 s instead of seconds
    rank, world = 0, 1
    solver = None
    if cfg.backend != "cpu":
        from .parallel.dist import init_distributed
        rank, world, local, device = init_distributed()
        if not torch.cuda.is_available() and cfg.backend != "nccl-halo":
            print("no CUDA device: use --backend cpu (oracle) or --backend nccl-halo under torchrun",
                  file=sys.stderr)
            return 1
    try:
        if cfg.backend not in ("cpu", "single", "resident", "plain") or world > 1:
            cfg.validate(world_size=world if world > 1 or cfg.backend == "nccl-halo" else None)
        from .api import make_solver
        solver = make_solver(cfg)
    except (ConfigError, ValueError) as e:
        if rank == 0:
            print(str(e), file=sys.stderr)
            print("Quitting...", file=sys.stderr)
        return 1

    from .utils import dump as D
    from .utils.timing import PhaseTimer
    timer = PhaseTimer(torch.cuda.current_device() if (cfg.backend != "cpu" and torch.cuda.is_available()) else None)
    style = a.preset or "grad1612_mpi"
    if rank == 0:
        _banner(cfg, world if world > 1 else cfg.px * cfg.py, style)
    if cfg.debug:
        from .parallel.decomp import Decomposition
        dec = Decomposition(cfg.nx, cfg.ny, cfg.px, cfg.py, cfg.uneven)
        # grad1612_mpi_heat.c:170-175; a missing neighbour prints as MPI_PROC_NULL, which is
        # -2 in Open MPI and in the shim under baseline/mpishim.  One process that emulates
        # P x Q tiles speaks for all of them.
        null = lambda q: -2 if q is None else q
        for t in (dec.tiles if world == 1 else [dec.tile(rank)]):
            nb = t.neighbours
            print(f"I am {t.rank} and my neighbors are North={null(nb['N'])}, South={null(nb['S'])}, "
                  f"East={null(nb['E'])}, West={null(nb['W'])} (Running on {os.uname().nodename})")
        if rank == 0 and cfg.backend != "cpu" and torch.cuda.is_available():
            # every visible device, as detailsGPU() does (grad1612_cuda_heat.cu:24-37)
            from .ops import kernels as K
            for dev in range(torch.cuda.device_count()):
                print(K.device_info(dev))

    with timer.phase("init"):
        if cfg.restart:
            grid = D.read_binary(cfg.restart, cfg.nx, cfg.ny)
            solver.load(grid)
        else:
            solver.initialize()

    def write(stem):
        if not dumps:
            return
        os.makedirs(cfg.out_dir, exist_ok=True)
        bin_path = os.path.join(cfg.out_dir, f"{stem}_binary.dat")
        txt_path = os.path.join(cfg.out_dir, f"{stem}.dat")
        if world > 1:
            # Every rank writes its own tile straight into the shared file at its place in the
            # nx x ny array — the MPI-IO subarray view of grad1612_mpi_heat.c:178-190,283-285 —
            # and rank 0 then converts the binary file to text exactly as the reference does
            # (:191-203,286-298).  Nothing is gathered: at 32768^2 a gather would funnel 4 GiB
            # through one rank and one PCIe link.
            import torch.distributed as dist
            keep = "binary" in dumps
            path = bin_path if keep else os.path.join(cfg.out_dir, f".{stem}_binary.tmp")
            if rank == 0:
                open(path, "ab").close()     # create without truncating (the reference opens WRONLY|CREATE)
            dist.barrier()
            for x0, y0, tile in solver.local_tiles():
                D.write_binary_tile(path, tile, cfg.nx, cfg.ny, x0, y0, truncate=False)
            dist.barrier()
            if rank == 0:
                for dname in dumps:
                    if dname in ("prtdat", "rowmajor"):
                        D.binary_to_text(path, txt_path, cfg.nx, cfg.ny, dname)
                if not keep:
                    os.remove(path)
            return
        field = solver.gather(as_numpy=True) if cfg.backend == "cpu" else solver.gather()
        for dname in dumps:
            if dname == "binary":
                D.write_binary(bin_path, field, truncate=False)
            elif dname == "prtdat":
                D.write_prtdat(txt_path, field)
            else:
                D.write_rowmajor(txt_path, field)

    if rank == 0 and dumps:
        print("Writing initial.dat ..." if style != "mpi_heat2Dn"
              else "Initializing grid and writing initial.dat file...")
    with timer.phase("dump_initial"):
        write("initial")
    if style == "mpi_heat2Dn" and rank == 0:
        # The master's assignment log (mpi_heat2Dn.c:113-114) and the workers' start line
        # (:173).  Workers are numbered from 1 there (task 0 is the idle master); NONE = 0.
        from .parallel.decomp import split
        nw = cfg.px * cfg.py
        for i, (off, rows) in enumerate(split(cfg.nx, cfg.px), start=1):
            left, right = (i - 1 if i > 1 else 0), (i + 1 if i < nw else 0)
            print(f"Sent to task {i}: rows= {rows} offset= {off} left= {left} right= {right}")
        for i in range(1, nw + 1):
            print(f"Task {i} received work. Beginning time steps...")
    with timer.phase("run"):
        res = solver.run()
    if rank == 0:
        if style == "mpi_heat2Dn":
            print(f"Elapsed time: {res.elapsed_s:e} sec")
            if dumps:
                print("Writing final.dat file and generating graph...")
        elif style == "grad1612_cuda":
            print(f"Elapsed time: {res.elapsed_s:e} sec")
        else:
            print(f"Exiting after {res.iterations} iterations\nElapsed time: {res.elapsed_s:e} sec")
            if dumps:
                print("Writing final.dat ...")
    with timer.phase("dump_final"):
        write("final")
    if rank == 0:
        rec = {"phases": timer.as_dict(),"config": cfg.to_dict(), "world": world, "steps_done": res.steps_done,
               "iterations": res.iterations, "converged": res.converged, "last_norm": res.last_norm,
               "seconds_max_over_ranks": res.elapsed_s, "glups": res.glups, "launches": res.launches,
               "checks": res.checks, "when": time.strftime("%Y-%m-%dT%H:%M:%S")}
        try:
            with open(os.path.join(os.path.dirname(os.path.dirname(__file__)), "MEASURED_PEAKS.json")) as f:
                hbm = json.load(f).get("hbm_gbs")
            if hbm and cfg.backend not in ("cpu",):
                rec["pct_of_k1_hbm_roofline"] = 100.0 * res.glups / (hbm / 8.0 * max(world, 1))
            if cfg.backend in ("fused", "single"):
                from .utils.perfmodel import roofline_fraction
                rec["roofline"] = roofline_fraction(res.glups, cfg.nx, cfg.ny, cfg.px, cfg.py, max(cfg.k, 1),
                                                    cfg.chunk_rows)
        except Exception:  # noqa: BLE001
            pass
        os.makedirs(cfg.out_dir, exist_ok=True)
        with open(os.path.join(cfg.out_dir, "result.json"), "w") as f:
            json.dump(rec, f, indent=1)
        print(f"{res.glups:.3f} GLUP/s")
    close = getattr(solver, "close", None)
    if close:
        close()
    if cfg.backend != "cpu":
        try:
            import torch.distributed as dist
            if dist.is_initialized():
                dist.barrier()
                dist.destroy_process_group()
        except Exception:  # noqa: BLE001
            pass
    return 0


if __name__ == "__main__":
    sys.exit(main())
