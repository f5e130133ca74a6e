"""CLI: reference banner lines, dump files, JSON record (cpu backend here; the GPU
run of the same entry point is in test_gpu_kernels.py::test_cli_gpu)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from heat2d_b200 import _native as N
from heat2d_b200.cli import build_parser, config_from_args
from heat2d_b200.models.presets import PRESETS, preset
from heat2d_b200.ops import kernels as K
from heat2d_b200.utils import dump as D

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class _Res:
    def __init__(self, rc, out, err):
        self.returncode, self.stdout, self.stderr = rc, out, err


def run_cli(*args):
    """Run the CLI in-process (the module entry point itself is covered by test_module_entry_point)."""
    import contextlib
    import io
    from heat2d_b200.cli import main
    out, err = io.StringIO(), io.StringIO()
    with contextlib.redirect_stdout(out), contextlib.redirect_stderr(err):
        rc = main(list(args))
    return _Res(rc, out.getvalue(), err.getvalue())


def test_module_entry_point(tmp_path):
    r = subprocess.run([sys.executable, "-m", "heat2d_b200", "--backend", "cpu", "--out-dir", str(tmp_path)],
                       cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "Problem size:10x10" in r.stdout


def test_presets_cover_the_four_programs():
    assert set(PRESETS) == {"mpi_heat2Dn", "grad1612_mpi", "grad1612_hybrid", "grad1612_cuda"}
    c = preset("grad1612_cuda")
    assert (c.nx, c.ny, c.steps, c.boundary, c.arith, c.init) == (640, 1024, 10000, "fixed", "strict-fma", "ref-uint32")
    h = preset("grad1612_hybrid")
    assert h.conv and h.threads == 4 and (h.px, h.py) == (1, 1)
    assert preset("mpi_heat2Dn").cx_float


def test_flags_override_preset():
    a = build_parser().parse_args(["--preset", "grad1612_mpi", "--nx", "80", "--ny", "64", "--conv", "on", "-k", "4"])
    c = config_from_args(a)
    assert (c.nx, c.ny, c.conv, c.k, c.boundary) == (80, 64, True, 4, "ghost-ring")


def test_cpu_run_grad1612_style(tmp_path):
    r = run_cli("--preset", "grad1612_mpi", "--backend", "cpu", "--px", "1", "--py", "1", "--out-dir", str(tmp_path))
    assert r.returncode == 0, r.stderr
    out = r.stdout
    assert "Problem size:10x10" in out and "Amount of iterations: 100" in out
    assert "Exiting after 100 iterations" in out and "Elapsed time: " in out and " sec" in out
    want = K.cpu_initial_grid(10, 10, "ref-int32")
    assert np.array_equal(D.read_binary(str(tmp_path / "initial_binary.dat"), 10, 10), want)
    K.cpu_run(want, 100, N.make_phys(0.1, 0.1, "ghost-ring", "strict"))
    assert np.array_equal(D.read_binary(str(tmp_path / "final_binary.dat"), 10, 10), want)
    lines = (tmp_path / "final.dat").read_text().splitlines()
    assert len(lines) == 10 and all(l.endswith(" ") for l in lines)
    rec = json.loads((tmp_path / "result.json").read_text())
    assert rec["steps_done"] == 100 and rec["config"]["boundary"] == "ghost-ring"


def test_cpu_run_given_program_style(tmp_path):
    r = run_cli("--preset", "mpi_heat2Dn", "--backend", "cpu", "--px", "1", "--out-dir", str(tmp_path))
    assert r.returncode == 0, r.stderr
    assert "Grid size: X= 10  Y= 10  Time steps= 100" in r.stdout
    txt = (tmp_path / "final.dat").read_text()
    assert txt.count("\n") == 10 and not txt.splitlines()[0].endswith(" ")


def test_convergence_and_bug_replay(tmp_path):
    r = run_cli("--preset", "grad1612_hybrid", "--backend", "cpu", "--steps", "100000", "--out-dir", str(tmp_path))
    assert r.returncode == 0, r.stderr
    rec = json.loads((tmp_path / "result.json").read_text())
    assert rec["converged"] and rec["iterations"] % 20 == 0
    # shipped predicate: tile height 10 is not a multiple of 20 -> the check never fires
    r = run_cli("--preset", "grad1612_hybrid", "--backend", "cpu", "--steps", "500", "--ref-conv-bug",
                "--out-dir", str(tmp_path))
    rec = json.loads((tmp_path / "result.json").read_text())
    assert not rec["converged"] and rec["steps_done"] == 500


def test_bad_process_grid_message():
    r = run_cli("--nx", "10", "--ny", "10", "--px", "3", "--py", "2", "--no-uneven", "--backend", "cpu")
    assert r.returncode == 1 and "is not an integer" in r.stderr and "Quitting" in r.stderr


def test_restart_continues_a_run(tmp_path):
    a, b = tmp_path / "a", tmp_path / "b"
    run_cli("--nx", "24", "--ny", "20", "--steps", "30", "--backend", "cpu", "--dump", "binary", "--out-dir", str(a))
    r = run_cli("--nx", "24", "--ny", "20", "--steps", "20", "--backend", "cpu", "--dump", "binary",
                "--restart", str(a / "final_binary.dat"), "--start-step", "30", "--out-dir", str(b))
    assert r.returncode == 0, r.stderr
    want = K.cpu_initial_grid(24, 20, "exact")
    K.cpu_run(want, 50, N.make_phys(0.1, 0.1, "fixed", "fp32"))
    assert np.array_equal(D.read_binary(str(b / "final_binary.dat"), 24, 20), want)


def test_given_program_worker_limits():
    # MINWORKER 3 / MAXWORKER 8 (mpi_heat2Dn.c:32-33): enforced for the preset on device backends
    a = build_parser().parse_args(["--preset", "mpi_heat2Dn", "--px", "2"])
    assert config_from_args(a).px == 2
    r = run_cli("--preset", "mpi_heat2Dn", "--px", "2", "--backend", "fused")
    assert r.returncode == 1 and "between 4 and 9" in r.stderr


def test_cpu_backend_runs_without_importing_torch(tmp_path):
    """`--backend cpu` needs neither PyTorch nor a GPU; importing torch would be most of the
    run time of a reference-size job."""
    import subprocess
    import sys
    code = ("import sys; from heat2d_b200 import cli; "
            f"rc = cli.main(['--preset', 'grad1612_hybrid', '--backend', 'cpu', '--debug', '--out-dir', r'{tmp_path}']); "
            "print('TORCH', 'torch' in sys.modules, rc)")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120,
                       env=dict(os.environ, PYTHONPATH=root))
    assert r.returncode == 0, r.stderr
    assert "TORCH False 0" in r.stdout
    assert (tmp_path / "final.dat").exists() and (tmp_path / "final_binary.dat").exists()
