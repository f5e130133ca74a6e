"""Worker for the multi-GPU tests: one rank per GPU (launched by torch.distributed.run)."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from heat2d_b200 import HeatConfig, _native as N  # noqa: E402
from heat2d_b200.ops import kernels as K  # noqa: E402
from heat2d_b200.parallel.dist import (DistributedHeat2DSolver, HaloExchangeSolver,  # noqa: E402
                                       init_distributed)


def main():
    spec = json.loads(sys.argv[1])
    rank, world, local, device = init_distributed("nccl")
    results = []
    ok_all = True
    for case in spec["cases"]:
        cfg = HeatConfig(**case["cfg"])
        cls = HaloExchangeSolver if cfg.backend == "nccl-halo" else DistributedHeat2DSolver
        s = cls(cfg, device)
        if case.get("random"):
            u0 = (np.random.default_rng(7).random((cfg.nx, cfg.ny), dtype=np.float32) * 1e4).astype(np.float32)
            if case.get("negative"):
                u0[cfg.nx // 3:cfg.nx // 2, :] *= np.float32(-1.0)   # the strict fast path must fall back there
            s.load(u0)
        else:
            u0 = K.cpu_initial_grid(cfg.nx, cfg.ny, cfg.init)
            s.initialize()
        if case.get("pipelined"):
            # column strips, end to end from / to pinned host memory, ranks in lockstep
            hin, col0, hout = s.host_strip()
            spec_t = s.decomp.tile(rank)
            c_lo, c_hi = spec_t.y0 - col0, spec_t.y0 + spec_t.cols + (hin.shape[1] - col0 - spec_t.cols)
            hin.copy_(torch.from_numpy(np.ascontiguousarray(u0[:, c_lo:c_hi])))
            res = s.solve_pipelined(hin, col0, hout)
            # a run() afterwards must find flags and epochs in a consistent state
            res2 = s.run(cfg.k)
            parts = [None] * world
            dist.all_gather_object(parts, (spec_t.y0, spec_t.cols, hout.numpy().copy()))
            out = torch.empty((cfg.nx, cfg.ny), dtype=torch.float32)
            for y0, cols, part in parts:
                out[:, y0:y0 + cols] = torch.from_numpy(part)
            after = s.gather()
            if rank == 0:
                ref2 = u0.copy()
                K.cpu_run(ref2, cfg.steps + cfg.k, N.make_phys(cfg.cx_eff, cfg.cy_eff, cfg.boundary, cfg.arith),
                          "none", 20, 0.1, threads=8)
                if not np.array_equal(after.numpy(), ref2) or res2.steps_done != cfg.steps + cfg.k:
                    ok_all = False
                    results.append({"name": case.get("name", "") + " (run after pipelined)", "ok": False})
        else:
            res = s.run()
            out = s.gather()
        if rank == 0:
            ref = u0.copy()
            done, last = K.cpu_run(ref, cfg.steps, N.make_phys(cfg.cx_eff, cfg.cy_eff, cfg.boundary, cfg.arith),
                                   cfg.norm if cfg.conv else "none", cfg.interval, cfg.eps, threads=8)
            same = bool(np.array_equal(out.numpy(), ref))
            ok = same and res.steps_done == done
            ok_all = ok_all and ok
            results.append({"name": case.get("name", ""), "ok": ok, "same": same, "steps_done": res.steps_done,
                            "oracle_done": done, "converged": res.converged, "glups": res.glups})
        s.close()
        dist.barrier()
    if rank == 0:
        print(json.dumps({"ok": ok_all, "world": world, "results": results}))
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok_all else 1)


if __name__ == "__main__":
    main()
