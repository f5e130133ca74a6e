"""Worker for the multi-process CPU tests (launched by torch.distributed.run)."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from heat2d_b200 import HeatConfig, _native as N  # noqa: E402
from heat2d_b200.ops import kernels as K  # noqa: E402
from heat2d_b200.parallel.dist import HaloExchangeSolver, init_distributed  # noqa: E402


def main():
    spec = json.loads(sys.argv[1])
    rank, world, _, device = init_distributed("gloo")
    cfg = HeatConfig(**spec["cfg"])
    s = HaloExchangeSolver(cfg, torch.device("cpu"))
    if spec.get("random"):
        u0 = (np.random.default_rng(7).random((cfg.nx, cfg.ny), dtype=np.float32) * 1e4).astype(np.float32)
        s.load(u0)
    else:
        u0 = K.cpu_initial_grid(cfg.nx, cfg.ny, cfg.init)
        s.initialize()
    res = s.run()
    out = s.gather()
    ok = True
    if rank == 0:
        ref = u0.copy()
        done, last = K.cpu_run(ref, cfg.steps, N.make_phys(cfg.cx_eff, cfg.cy_eff, cfg.boundary, cfg.arith),
                               cfg.norm if cfg.conv else "none", cfg.interval, cfg.eps)
        ok = bool(np.array_equal(out.numpy(), ref)) and res.steps_done == done
        print(json.dumps({"ok": ok, "steps_done": res.steps_done, "oracle_done": done,
                          "converged": res.converged, "world": world}))
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
