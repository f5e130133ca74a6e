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


def test_cli_dumps_are_written_per_rank_and_equal_the_single_rank_files(tmp_path):
    """Four ranks write their tiles straight into the shared binary file (the MPI-IO subarray
    write of grad1612_mpi_heat.c:178-190), rank 0 converts it to text: the bytes must equal the
    files of a one-process run."""
    ROOT = os.path.dirname(HERE)
    common = ["--nx", "50", "--ny", "70", "--steps", "12", "--arith", "strict", "--boundary", "ghost-ring",
              "--dump", "binary,rowmajor"]
    env = dict(os.environ, OMP_NUM_THREADS="1", CUDA_VISIBLE_DEVICES="")
    one, four = tmp_path / "one", tmp_path / "four"
    r = subprocess.run([sys.executable, "-m", "heat2d_b200", *common, "--backend", "cpu", "--out-dir", str(one)],
                       capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    four.mkdir()
    # a stale, longer file must keep its tail (the reference opens without truncating) but every
    # cell of the grid must be overwritten
    (four / "final_binary.dat").write_bytes(b"\xff" * (50 * 70 * 4 + 16))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=4",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), "-m", "heat2d_b200", *common,
           "--backend", "nccl-halo", "-k", "1", "--px", "2", "--py", "2", "--out-dir", str(four)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    for name in ("initial.dat", "final.dat", "initial_binary.dat"):
        assert (one / name).read_bytes() == (four / name).read_bytes(), name
    got = (four / "final_binary.dat").read_bytes()
    assert got[:50 * 70 * 4] == (one / "final_binary.dat").read_bytes() and got[50 * 70 * 4:] == b"\xff" * 16
