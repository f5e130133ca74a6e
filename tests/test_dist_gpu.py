"""One process per GPU (needs >= 2 GPUs): the fused peer-memory halo exchange and
the cross-GPU convergence finish against the CPU oracle, bit for bit."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _ngpu():
    try:
        return torch.cuda.device_count()
    except Exception:
        return 0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(nproc, spec, timeout=600, env_extra=None):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(HERE, "_dist_gpu_worker.py"), json.dumps(spec)]
    env = dict(os.environ)
    if env_extra:
        env.update(env_extra)
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    return json.loads(line)


def _cases(px, py):
    base = dict(nx=1000, ny=1400, px=px, py=py, k=8)
    return [
        {"name": "fused-k8", "random": True, "cfg": dict(base, steps=83, backend="fused")},
        {"name": "fused-ring-k5", "random": True,
         "cfg": dict(base, steps=42, backend="fused", k=5, boundary="ghost-ring")},
        {"name": "fused-strict", "random": True, "cfg": dict(base, steps=24, backend="fused", arith="strict")},
        {"name": "fused-conv", "cfg": dict(nx=48, ny=64, px=px, py=py, k=8, steps=100000, backend="fused",
                                           boundary="ghost-ring", conv=True, interval=20, eps=0.1)},
        {"name": "nccl-halo", "random": True, "cfg": dict(base, steps=9, backend="nccl-halo", k=1)},
        # the fair baseline: same temporal-blocked kernel, k-wide bands over NCCL once per launch
        {"name": "nccl-halo-k8", "random": True, "cfg": dict(base, steps=43, backend="nccl-halo", k=8)},
        {"name": "fused-k8-ab4", "random": True, "cfg": dict(base, steps=41, backend="fused", layout="ab4")},
        {"name": "fused-strict-fma-mixed-sign", "random": True, "negative": True,
         "cfg": dict(base, steps=24, backend="fused", arith="strict-fma")},
        # k = 4: what best_k picks for strict arithmetic on large tiles (bench.py's headline)
        {"name": "fused-strict-fma-k4", "random": True, "cfg": dict(base, steps=42, backend="fused", k=4,
                                                                     arith="strict-fma")},
    ]


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("px,py", [(2, 1), (1, 2)])
def test_two_gpus(px, py):
    out = _run(2, {"cases": _cases(px, py)})
    assert out["ok"], out


@pytest.mark.skipif(_ngpu() < 4, reason="needs 4 GPUs")
def test_four_gpus():
    out = _run(4, {"cases": _cases(2, 2)})
    assert out["ok"], out


@pytest.mark.skipif(_ngpu() < 8, reason="needs 8 GPUs")
def test_eight_gpus():
    out = _run(8, {"cases": _cases(2, 4)})
    assert out["ok"], out


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
def test_flag_protocol_under_injected_delays():
    """Race detection for the fused halo exchange: every edge task sleeps a pseudo-random
    time (up to ~64 us) before it touches the flag protocol, on every rank differently;
    the result must still be bit-identical to the oracle (a missing RAW/WAR edge in the
    protocol shows up as a mismatch)."""
    n = min(_ngpu(), 4)
    px, py = (2, 2) if n >= 4 else (1, 2)
    cases = [{"name": "delay", "random": True,
              "cfg": dict(nx=768, ny=1024, px=px, py=py, k=8, steps=120, backend="fused", chunk_rows=32)}]
    out = _run(px * py, {"cases": cases}, env_extra={"H2D_DEBUG_DELAY": "1"})
    assert out["ok"], out


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("delay", [False, True])
def test_pipelined_solve_on_column_strips(delay):
    """End to end on a 1 x P grid: upload / window launches with W-E halo exchange / download,
    all ranks cutting the same row blocks; then an ordinary run() on the same solver."""
    n = min(_ngpu(), 4)
    cases = [
        {"name": "pipelined-k8", "random": True, "pipelined": True,
         "cfg": dict(nx=3000, ny=256 * n, px=1, py=n, k=8, steps=83, backend="fused", chunk_rows=64)},
        {"name": "pipelined-k5-ring", "random": True, "pipelined": True,
         "cfg": dict(nx=1500, ny=192 * n, px=1, py=n, k=5, steps=41, backend="fused", boundary="ghost-ring")},
        # the last strip owns 4 columns (cols = 2 * 112 + 4): the strip before it depends on the E
        # ghost columns and must take part in the flag protocol as well
        {"name": "pipelined-k8-narrow-last-strip", "random": True, "pipelined": True,
         "cfg": dict(nx=2000, ny=228 * n, px=1, py=n, k=8, steps=50, backend="fused", chunk_rows=48, layout="v4")},
        # the sweep depth of bench.py's end-to-end run on large tiles (strict arithmetic, k = 4)
        {"name": "pipelined-k4-strict-fma", "random": True, "pipelined": True,
         "cfg": dict(nx=2500, ny=256 * n, px=1, py=n, k=4, steps=20, backend="fused", arith="strict-fma")},
    ]
    out = _run(n, {"cases": cases}, env_extra={"H2D_DEBUG_DELAY": "1"} if delay else None)
    assert out["ok"], out
