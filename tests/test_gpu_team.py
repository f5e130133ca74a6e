"""The cluster ("team") launch of the temporal-blocked kernel (csrc/tk_team.cuh,
csrc/stencil_team.cu): adjacent strips hand each other their edge columns through shared
memory / distributed shared memory instead of recomputing them.  Bit for bit against the CPU
oracle; the reference has one kernel and no tests (grad1612_cuda_heat.cu:55-62)."""
import numpy as np
import pytest
import torch

from heat2d_b200 import _native as N
from heat2d_b200.config import HeatConfig
from heat2d_b200.models.heat import Heat2DSolver, tensor_from_ptr
from heat2d_b200.ops import kernels as K

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def oracle(u0, steps, boundary, arith):
    u = u0.copy()
    K.cpu_run(u, steps, N.make_phys(0.1, 0.1, boundary, arith), threads=4)
    return u


def run_team(u0, k, launches, cluster, boundary, arith, chunk_rows=0, cx=0.1, cy=0.1):
    nx, ny = u0.shape
    cfg = HeatConfig(nx=nx, ny=ny, steps=k * launches, k=k, boundary=boundary, arith=arith, layout="v4", cx=cx, cy=cy)
    s = Heat2DSolver(cfg, DEV).load(u0)
    plan = s.engines[0].plan
    par = s.parity
    for _ in range(launches):
        par = plan.launch_team(par, k, cluster, chunk_rows)
    torch.cuda.synchronize()
    out = s.fields[0].owned(par).cpu().numpy()
    fb = tensor_from_ptr(plan.team_fallback_ptr, (1,), torch.int32, torch.device(DEV)).item()
    s.close()
    return out, fb


@pytest.mark.parametrize("cluster", [1, 2, 4])
@pytest.mark.parametrize("k", [5, 6, 7, 8])
@pytest.mark.parametrize("arith", ["fp32", "strict-fma"])
def test_team_launch_matches_the_oracle(cluster, k, arith):
    # wide enough for two teams of the largest cluster (the last one shifted), several chunks
    nx, ny = 131, 2 * (120 * 4 * cluster) + 77
    u0 = (np.random.default_rng(k + cluster).random((nx, ny), dtype=np.float32) * 1e4).astype(np.float32)
    for boundary in ("fixed", "ghost-ring"):
        got, fb = run_team(u0, k, 3, cluster, boundary, arith, chunk_rows=40)
        assert fb == 0
        assert np.array_equal(got, oracle(u0, 3 * k, boundary, arith)), (boundary,)


@pytest.mark.parametrize("cluster", [1, 2])
def test_team_launch_on_a_large_interior(cluster):
    # tasks that touch no boundary run the non-frozen body; automatic chunk height
    nx, ny = 2100, 4000
    u0 = (np.random.default_rng(3).random((nx, ny), dtype=np.float32) * 100).astype(np.float32)
    got, _ = run_team(u0, 8, 2, cluster, "fixed", "fp32")
    assert np.array_equal(got, oracle(u0, 16, "fixed", "fp32"))


def test_team_launch_counts_data_outside_the_fast_strict_domain():
    nx, ny = 64, 1100
    u0 = (np.random.default_rng(4).random((nx, ny), dtype=np.float32) * 100).astype(np.float32)
    u0[20, 500] = np.float32(-3.0)
    _, fb = run_team(u0, 8, 1, 2, "fixed", "strict-fma")
    assert fb > 0


@pytest.mark.parametrize("arith", ["strict", "strict-fma"])
def test_one_cta_teams_redo_tasks_outside_the_fast_strict_domain(arith):
    # negative, infinite and NaN cells: the teams that see them run again with the plain evaluation
    nx, ny = 150, 1500
    u0 = (np.random.default_rng(5).random((nx, ny), dtype=np.float32) * 100).astype(np.float32)
    u0[20:30, 500:520] *= np.float32(-1.0)
    u0[100, 1200] = np.float32("inf")
    got, fb = run_team(u0, 8, 2, 1, "fixed", arith, chunk_rows=48)
    want = oracle(u0, 16, "fixed", arith)
    assert fb > 0
    gn, wn = np.isnan(got), np.isnan(want)
    assert wn.any() and np.array_equal(gn, wn)
    assert np.array_equal(got.view(np.uint32)[~gn], want.view(np.uint32)[~wn])


@pytest.mark.parametrize("arith", ["fp32", "strict-fma"])
@pytest.mark.parametrize("conv", [False, True])
def test_solver_with_team_launches(arith, conv):
    # HeatConfig.team = "on": the k >= 5 sweeps without a norm are cluster launches, the checked
    # sweeps and the k < 5 remainder are not; same schedule, same bits, same stopping step
    nx, ny = 300, 1000
    u0 = K.cpu_initial_grid(nx, ny, "exact")
    cfg = HeatConfig(nx=nx, ny=ny, steps=131, k=8, arith=arith, team="on", conv=conv, interval=20, eps=1e-3,
                     boundary="ghost-ring")
    s = Heat2DSolver(cfg, DEV).initialize()
    res = s.run()
    ref = u0.copy()
    done, _ = K.cpu_run(ref, 131, N.make_phys(0.1, 0.1, "ghost-ring", arith), "l2sq" if conv else "none", 20, 1e-3)
    assert res.steps_done == done
    assert np.array_equal(s.gather().numpy(), ref)
    # and the team kernels really ran
    assert s.engines[0].plan._desc.team == 1
    s.close()


def test_team_launch_rejects_what_it_cannot_do():
    s = Heat2DSolver(HeatConfig(nx=64, ny=300, steps=8, k=8, layout="v4"), DEV).initialize()
    with pytest.raises(RuntimeError, match="narrower than one team"):
        s.engines[0].plan.launch_team(0, 8, 2)
    with pytest.raises(RuntimeError, match="outside"):
        s.engines[0].plan.launch_team(0, 4, 1)
    s.close()
