"""Multi-process host-side logic on CPU (gloo): neighbour mapping, halo
pack/unpack order, uneven tiles, distributed convergence — the algorithm of
grad1612_mpi_heat.c:231-275 on the HaloExchangeSolver baseline."""
import json
import os
import socket
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(nproc, spec):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(HERE, "_dist_worker.py"), json.dumps(spec)]
    env = dict(os.environ, OMP_NUM_THREADS="1", CUDA_VISIBLE_DEVICES="")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    return json.loads(line)


@pytest.mark.parametrize("px,py", [(2, 2), (4, 1), (1, 3)])
def test_halo_exchange_equals_oracle(px, py):
    spec = {"cfg": dict(nx=37, ny=29, steps=30, px=px, py=py, backend="nccl-halo", arith="strict",
                        boundary="ghost-ring", k=1), "random": True}
    out = _run(px * py, spec)
    assert out["ok"] and out["world"] == px * py


def test_distributed_convergence_stops_at_oracle_step():
    spec = {"cfg": dict(nx=16, ny=16, steps=100000, px=2, py=2, backend="nccl-halo", arith="strict",
                        boundary="ghost-ring", conv=True, interval=20, eps=0.1, k=1)}
    out = _run(4, spec)
    assert out["ok"] and out["converged"] and out["steps_done"] == out["oracle_done"]
