"""Throughput bookkeeping: `glups` of a run counts the updates of THAT call only."""
import pytest

from heat2d_b200.config import HeatConfig
from heat2d_b200.models.cpu import CpuSolver
from heat2d_b200.models.heat import RunResult


def test_glups_uses_the_steps_of_this_call():
    r = RunResult(steps_done=1024, iterations=1024, converged=False, last_norm=float("nan"), elapsed_s=0.5,
                  launches=125, checks=0, nx=1000, ny=2000, steps_run=1000)
    assert r.glups == pytest.approx(1000 * 2000 * 1000 / 0.5 / 1e9)
    # records written before the field existed fall back to steps_done
    old = RunResult(1024, 1024, False, float("nan"), 0.5, 125, 0, 1000, 2000)
    assert old.glups == pytest.approx(1000 * 2000 * 1024 / 0.5 / 1e9)
    assert RunResult(8, 8, False, 0.0, 0.0, 1, 0, 10, 10, 8).glups != RunResult(8, 8, False, 0.0, 0.0, 1, 0, 10, 10, 8).glups  # nan


def test_cpu_solver_repeated_runs_and_restart_offset():
    s = CpuSolver(HeatConfig(nx=40, ny=30, steps=10, backend="cpu", start_step=100)).initialize()
    a = s.run()
    b = s.run(5)
    assert (a.steps_done, a.steps_run) == (110, 10)
    assert (b.steps_done, b.steps_run) == (115, 5)
    assert b.glups == pytest.approx(40 * 30 * 5 / b.elapsed_s / 1e9)
