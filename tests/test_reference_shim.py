"""The UNMODIFIED reference MPI programs, built on the single-node MPI shim, against
this framework's oracle — byte-identical dumps (BASELINE config 1 and SURVEY 4 "Oracle").

Needs the reference sources (staged in baseline/_ref/src or present in /root/reference)
and gcc; skipped otherwise.
"""
import os
import re
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline"))
import ref_build as rb  # noqa: E402

from heat2d_b200 import _native as N  # noqa: E402
from heat2d_b200.models.presets import preset  # noqa: E402
from heat2d_b200.models.cpu import CpuSolver  # noqa: E402
from heat2d_b200.utils import dump as D  # noqa: E402


def _have_reference():
    try:
        if not rb.staged():
            rb.stage()
        rb.build_shim()
        return True
    except Exception:
        return False


pytestmark = pytest.mark.skipif(not _have_reference(), reason="reference sources or gcc not available")


def _ours(name, tmp, **over):
    cfg = preset(name, backend="cpu", **over)
    s = CpuSolver(cfg).initialize()
    init = s.u.copy()
    res = s.run()
    return cfg, init, s.u, res


def test_config1_given_program_640x640_on_4_ranks_is_byte_identical(tmp_path):
    # BASELINE.json config 1: mpi_heat2Dn.c-equivalent, 640x640, 100 steps, mpiexec -n 4
    exe = rb.build_mpi("mpi_heat2Dn", {"NXPROB": 640, "NYPROB": 640, "STEPS": 100}, "-O2")
    out, secs = rb.run_mpi(exe, 4, tmp_path)
    assert "Starting mpi_heat2D with 3 worker tasks." in out and secs is not None
    # the strip partition printed by the reference == our decomposition (214/213/213 rows)
    rows = [int(x) for x in re.findall(r"rows= (\d+)", out)]
    assert rows == [214, 213, 213]
    cfg, init, final, res = _ours("mpi_heat2Dn", tmp_path, nx=640, ny=640, steps=100, px=3)
    D.write_prtdat(str(tmp_path / "ours_initial.dat"), init)
    D.write_prtdat(str(tmp_path / "ours_final.dat"), final)
    assert (tmp_path / "initial.dat").read_bytes() == (tmp_path / "ours_initial.dat").read_bytes()
    assert (tmp_path / "final.dat").read_bytes() == (tmp_path / "ours_final.dat").read_bytes()


@pytest.mark.parametrize("workers", [3, 5, 8])
def test_given_program_any_worker_count(tmp_path, workers):
    exe = rb.build_mpi("mpi_heat2Dn", {"NXPROB": 50, "NYPROB": 37, "STEPS": 60}, "-O2")
    rb.run_mpi(exe, workers + 1, tmp_path)
    _, _, final, _ = _ours("mpi_heat2Dn", tmp_path, nx=50, ny=37, steps=60, px=workers)
    D.write_prtdat(str(tmp_path / "ours.dat"), final)
    assert (tmp_path / "final.dat").read_bytes() == (tmp_path / "ours.dat").read_bytes()


def test_given_program_rejects_bad_worker_count(tmp_path):
    exe = rb.build_mpi("mpi_heat2Dn", {"NXPROB": 10, "NYPROB": 10, "STEPS": 10}, "-O2")
    with pytest.raises(RuntimeError, match="number of tasks must be between 4 and 9"):
        rb.run_mpi(exe, 2, tmp_path)


def test_grad1612_mpi_single_rank_is_byte_identical(tmp_path):
    exe = rb.build_mpi("grad1612_mpi", {"NXPROB": 80, "NYPROB": 64, "STEPS": 100, "GRIDX": 1, "GRIDY": 1}, "-O2")
    out, secs = rb.run_mpi(exe, 1, tmp_path)
    assert "Exiting after 100 iterations" in out
    _, init, final, _ = _ours("grad1612_mpi", tmp_path, nx=80, ny=64, steps=100, px=1, py=1)
    assert np.array_equal(D.read_binary(str(tmp_path / "initial_binary.dat"), 80, 64), init)
    assert np.array_equal(D.read_binary(str(tmp_path / "final_binary.dat"), 80, 64), final)
    D.write_rowmajor(str(tmp_path / "ours.dat"), final)
    assert (tmp_path / "final.dat").read_bytes() == (tmp_path / "ours.dat").read_bytes()


def test_grad1612_mpi_four_ranks_initial_dump_and_the_final_dump_defect(tmp_path):
    # The subarray file view places every rank's tile correctly in initial_binary.dat ...
    exe = rb.build_mpi("grad1612_mpi", {"NXPROB": 12, "NYPROB": 16, "STEPS": 30, "GRIDX": 2, "GRIDY": 2}, "-O2")
    rb.run_mpi(exe, 4, tmp_path)
    _, init, final, _ = _ours("grad1612_mpi", tmp_path, nx=12, ny=16, steps=30, px=2, py=2)
    assert np.array_equal(D.read_binary(str(tmp_path / "initial_binary.dat"), 12, 16), init)
    # ... but the reference never calls MPI_File_set_view before the FINAL write
    # (grad1612_mpi_heat.c:283-285), so every rank writes its tile at offset 0 of
    # final_binary.dat: with more than one rank the reference's final dump holds
    # one tile, not the grid.  Recorded here so nobody tries to match it.
    assert os.path.getsize(tmp_path / "final_binary.dat") == 6 * 8 * 4


def test_grad1612_mpi_rejects_wrong_process_count(tmp_path):
    exe = rb.build_mpi("grad1612_mpi", {"NXPROB": 12, "NYPROB": 16, "STEPS": 30, "GRIDX": 2, "GRIDY": 2}, "-O2")
    with pytest.raises(RuntimeError, match="number of tasks must be equal to 4"):
        rb.run_mpi(exe, 2, tmp_path)


def test_grad1612_convergence_with_the_shipped_predicate(tmp_path):
    # 20x20 on one rank: the stale loop variable equals the tile height 20, 20 % INTERVAL == 0,
    # so the shipped code checks on EVERY step.  --ref-conv-bug replays that.
    exe = rb.build_mpi("grad1612_mpi", {"NXPROB": 20, "NYPROB": 20, "STEPS": 100000, "GRIDX": 1, "GRIDY": 1,
                                        "CONVERGENCE": 1}, "-O2")
    out, _ = rb.run_mpi(exe, 1, tmp_path)
    it = int(re.search(r"Exiting after (\d+) iterations", out).group(1))
    cfg = preset("grad1612_mpi", backend="cpu", nx=20, ny=20, steps=100000, px=1, py=1, conv=True, interval=1)
    s = CpuSolver(cfg).initialize()
    res = s.run()
    assert res.converged and abs(res.steps_done - it) <= 1   # float vs double accumulation of the norm
    if res.steps_done == it:
        assert np.array_equal(D.read_binary(str(tmp_path / "final_binary.dat"), 20, 20), s.u)


def test_grad1612_hybrid_runs_on_the_shim_and_matches(tmp_path):
    # MPI + OpenMP program, one rank, one thread (the shipped nested pragmas make >1 thread
    # redo the whole sweep per thread: same values, 4x the time — Report.pdf p24)
    exe = rb.build_mpi("grad1612_hybrid", {"NXPROB": 24, "NYPROB": 20, "STEPS": 60, "NUMTHREADS": 1,
                                           "CONVERGENCE": 0}, "-O2")
    out, secs = rb.run_mpi(exe, 1, tmp_path)
    assert secs is not None
    cfg = preset("grad1612_hybrid", backend="cpu", nx=24, ny=20, steps=60, conv=False, threads=1)
    s = CpuSolver(cfg).initialize()
    s.run()
    assert np.array_equal(D.read_binary(str(tmp_path / "final_binary.dat"), 24, 20), s.u)
