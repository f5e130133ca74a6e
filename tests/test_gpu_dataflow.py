"""Persistent dataflow launch (h2d_plan_run_dataflow): one launch runs many k-step sweeps,
tasks of sweep e+1 start as soon as their predecessors of sweep e are done.  Results must be
bit-identical to the launch-per-sweep path and to the CPU oracle."""
import numpy as np
import pytest
import torch

from heat2d_b200 import _native as N
from heat2d_b200.config import HeatConfig
from heat2d_b200.models.heat import Heat2DSolver
from heat2d_b200.ops import kernels as K

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def rand_grid(nx, ny, seed=0):
    return (np.random.default_rng(seed).random((nx, ny), dtype=np.float32) * 1e4).astype(np.float32)


def run(u0, dataflow, **kw):
    nx, ny = u0.shape
    s = Heat2DSolver(HeatConfig(nx=nx, ny=ny, dataflow=dataflow, **kw), DEV).load(u0)
    res = s.run()
    out = s.gather().numpy().copy()
    s.close()
    return out, res


@pytest.mark.parametrize("nx,ny,k,chunk,sweeps,boundary", [
    (10, 10, 8, 0, 2, "fixed"),              # one task
    (150, 333, 8, 32, 5, "fixed"),           # 5 chunks x 3 strips, partial last strip
    (150, 333, 3, 16, 7, "ghost-ring"),      # H = 4: other strip width; every task is a boundary task
    (641, 130, 4, 8, 9, "fixed"),            # many short chunks, two strips (the shifted last one overlaps the first)
    (1030, 2050, 5, 64, 9, "fixed"),
    (2048, 4096, 8, 32, 6, "fixed"),         # 2368 tasks: more than the resident warps of the GPU
])
def test_dataflow_equals_one_launch_per_sweep(nx, ny, k, chunk, sweeps, boundary):
    u0 = rand_grid(nx, ny, k + sweeps)
    kw = dict(steps=k * sweeps, k=k, chunk_rows=chunk, boundary=boundary)
    a, ra = run(u0, False, **kw)
    b, rb = run(u0, True, **kw)
    assert ra.launches == sweeps and rb.launches == 1
    assert rb.steps_done == k * sweeps
    assert np.array_equal(a, b)


@pytest.mark.parametrize("arith", ["fp32", "strict", "strict-fma"])
def test_dataflow_matches_the_oracle(arith):
    nx, ny = 200, 300
    u0 = rand_grid(nx, ny, 5)
    got, res = run(u0, True, steps=8 * 4 + 3, k=8, chunk_rows=24, arith=arith)
    # 35 steps at k = 8: five sweeps of 7 (balanced schedule), all in one dataflow launch
    assert res.launches == 1
    want = u0.copy()
    K.cpu_run(want, 35, N.make_phys(0.1, 0.1, "fixed", arith))
    assert np.array_equal(got, want)


def test_dataflow_general_coefficients_and_repeated_runs():
    nx, ny = 333, 555
    u0 = rand_grid(nx, ny, 9)
    cfg = dict(nx=nx, ny=ny, steps=24, k=8, chunk_rows=40, cx=0.07, cy=0.13)
    s = Heat2DSolver(HeatConfig(dataflow=True, **cfg), DEV).load(u0)
    r = Heat2DSolver(HeatConfig(dataflow=False, **cfg), DEV).load(u0)
    for _ in range(3):   # the per-task counters keep counting across calls
        s.run()
        r.run()
        assert np.array_equal(s.gather().numpy(), r.gather().numpy())
    assert s.steps_done == 72


def test_dataflow_native_entry_point_checks_arguments():
    nx, ny = 64, 256
    s = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=8, k=8), DEV).initialize()
    plan = s.engines[0].plan
    assert plan.run_dataflow(0, 8, 0) == 0           # nothing to do
    with pytest.raises(N.NativeError):
        plan.run_dataflow(2, 8, 1)
    with pytest.raises(N.NativeError):
        plan.run_dataflow(0, 9, 1)
    with pytest.raises(N.NativeError):
        plan.run_dataflow(0, 8, -1)
    # virtual tiles have neighbours: the dataflow variant refuses them
    v = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=8, k=8, px=2, py=1), DEV).initialize()
    with pytest.raises(N.NativeError):
        v.engines[0].plan.run_dataflow(0, 8, 2)
    # ... and the solver silently keeps the launch-per-sweep path for them
    u0 = rand_grid(nx, ny, 1)
    a = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=24, k=8, px=2, py=1, dataflow=True), DEV).load(u0)
    a.run()
    b = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=24, k=8), DEV).load(u0)
    b.run()
    assert np.array_equal(a.gather().numpy(), b.gather().numpy())
    torch.cuda.synchronize()
