"""The lane layouts of the fp32 kernels (v4 / ab2 / ab4, csrc/tk_task.cuh), the fast
evaluation of the strict modes with its fallback, and the failure paths — all against the
CPU oracle, bit for bit.  The reference has one kernel and no tests
(grad1612_cuda_heat.cu:55-62)."""
import numpy as np
import pytest
import torch

from heat2d_b200 import _native as N
from heat2d_b200.config import HeatConfig
from heat2d_b200.models.heat import Heat2DSolver
from heat2d_b200.ops import kernels as K

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
LAYOUTS = ["v4", "ab2", "ab4"]


def oracle(u0, steps, boundary="fixed", arith="fp32", cx=0.1, cy=0.1):
    u = u0.copy()
    K.cpu_run(u, steps, N.make_phys(cx, cy, boundary, arith), threads=4)
    return u


def same_bits_or_nan(got, want):
    """Bit-identical, except that a NaN only has to be a NaN: x86 and the GPU disagree on the sign
    and payload of generated NaNs (inf - inf is 0xffc00000 on x86, 0x7fffffff on sm_100)."""
    gn, wn = np.isnan(got), np.isnan(want)
    return np.array_equal(gn, wn) and np.array_equal(got.view(np.uint32)[~gn], want.view(np.uint32)[~wn])


def rand_grid(nx, ny, seed=0, scale=1e4):
    return (np.random.default_rng(seed).random((nx, ny), dtype=np.float32) * scale).astype(np.float32)


@pytest.mark.parametrize("layout", LAYOUTS)
@pytest.mark.parametrize("k", [1, 2, 3, 4, 5, 6, 7, 8])
@pytest.mark.parametrize("boundary", ["fixed", "ghost-ring"])
def test_every_layout_every_k_bit_exact(layout, k, boundary):
    nx, ny = 150, 613      # several strips in every layout, partial last strip, odd sizes
    u0 = rand_grid(nx, ny, k)
    cfg = HeatConfig(nx=nx, ny=ny, steps=2 * k + 1, k=k, boundary=boundary, chunk_rows=32, layout=layout)
    s = Heat2DSolver(cfg, DEV).load(u0)
    assert s.engines[0].plan.layout == layout
    res = s.run()
    assert res.steps_done == 2 * k + 1
    assert np.array_equal(s.gather().numpy(), oracle(u0, 2 * k + 1, boundary))


@pytest.mark.parametrize("layout", ["ab2", "ab4"])
@pytest.mark.parametrize("shape", [(10, 10), (80, 64), (640, 512), (1030, 2050), (97, 240), (64, 241)])
def test_packed_layouts_on_reference_sizes(layout, shape):
    nx, ny = shape
    u0 = K.cpu_initial_grid(nx, ny, "exact")
    s = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=100, k=8, layout=layout), DEV).initialize()
    s.run()
    assert np.array_equal(s.gather().numpy(), oracle(u0, 100))


@pytest.mark.parametrize("layout", ["ab2", "ab4"])
def test_packed_layouts_general_coefficients(layout):
    nx, ny = 200, 700
    u0 = rand_grid(nx, ny, 3)
    s = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=17, k=8, cx=0.07, cy=0.13, layout=layout), DEV).load(u0)
    s.run()
    assert np.array_equal(s.gather().numpy(), oracle(u0, 17, cx=0.07, cy=0.13))


@pytest.mark.parametrize("layout", ["ab2", "ab4"])
@pytest.mark.parametrize("px,py", [(2, 2), (3, 1), (1, 3), (2, 4)])
@pytest.mark.parametrize("boundary", ["fixed", "ghost-ring"])
def test_packed_layouts_virtual_tiles(layout, px, py, boundary):
    # the halo-forwarding code of the packed task body: row bands, batched column bands
    # (tiles >= 64 columns wide, a multiple of 4) and the row-by-row fallback (odd widths)
    for nx, ny in ((203, 317), (256, 1024)):
        u0 = rand_grid(nx, ny, 9)
        cfg = HeatConfig(nx=nx, ny=ny, steps=41, k=8, px=px, py=py, boundary=boundary, chunk_rows=16, layout=layout)
        s = Heat2DSolver(cfg, DEV).load(u0)
        res = s.run()
        assert res.steps_done == 41
        assert np.array_equal(s.gather().numpy(), oracle(u0, 41, boundary)), (nx, ny)


@pytest.mark.parametrize("layout", ["ab2", "ab4"])
@pytest.mark.parametrize("px,py", [(1, 1), (2, 2)])
def test_packed_layouts_convergence(layout, px, py):
    nx, ny = 48, 320
    cfg = HeatConfig(nx=nx, ny=ny, steps=100000, k=8, px=px, py=py, boundary="ghost-ring", conv=True,
                     interval=20, eps=0.1, layout=layout)
    s = Heat2DSolver(cfg, DEV).initialize()
    res = s.run()
    ref = K.cpu_initial_grid(nx, ny, "exact")
    done, _ = K.cpu_run(ref, 100000, N.make_phys(0.1, 0.1, "ghost-ring", "fp32"), "l2sq", 20, 0.1)
    assert res.converged and res.iterations == done
    assert np.array_equal(s.gather().numpy(), ref)


@pytest.mark.parametrize("layout", ["ab2", "ab4"])
def test_packed_layouts_pipelined_solve(layout):
    nx, ny = 1030, 2050
    u0 = rand_grid(nx, ny, 4)
    hin = torch.from_numpy(u0).pin_memory()
    hout = torch.empty((nx, ny), dtype=torch.float32).pin_memory()
    s = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=43, k=8, layout=layout), DEV)
    s.solve_pipelined(hin, hout, 43)
    assert np.array_equal(hout.numpy(), oracle(u0, 43))


# ------------------------------------------------------------------ strict arithmetic
@pytest.mark.parametrize("arith", ["strict", "strict-fma"])
@pytest.mark.parametrize("k", [1, 3, 8])
def test_strict_fast_path_on_ordinary_data(arith, k):
    nx, ny = 300, 500
    u0 = rand_grid(nx, ny, 21)
    u0[:, 100:140] = 0.0                                  # zeros are inside the fast path's domain
    u0[50:60, :] *= np.float32(1e-42)                     # and so are subnormals
    s = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=3 * k + 1, k=k, arith=arith, chunk_rows=32), DEV).load(u0)
    s.run()
    assert np.array_equal(s.gather().numpy(), oracle(u0, 3 * k + 1, "fixed", arith))


@pytest.mark.parametrize("arith", ["strict", "strict-fma"])
def test_strict_fallback_on_negative_and_non_finite_data(arith):
    # every task sees a negative value: launch 1 retries task by task, the later launches go
    # straight to the plain evaluation (previous launch's counter); NaN / Inf must propagate
    nx, ny = 200, 400
    u0 = rand_grid(nx, ny, 22) - np.float32(5e3)
    u0[17, 33] = np.float32("inf")
    u0[150, 390] = np.float32("nan")
    u0[90, 200] = np.float32(3e38)
    s = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=40, k=8, arith=arith, chunk_rows=32), DEV).load(u0)
    s.run()
    got, want = s.gather().numpy(), oracle(u0, 40, "fixed", arith)
    assert np.isnan(want).any()
    assert same_bits_or_nan(got, want)


@pytest.mark.parametrize("px,py", [(2, 2), (1, 3)])
def test_strict_mixed_data_on_virtual_tiles(px, py):
    # some tasks fall back, most do not; fallback tasks re-send their halo bands
    nx, ny = 256, 640
    u0 = rand_grid(nx, ny, 23)
    u0[120:136, 300:340] *= np.float32(-1.0)
    cfg = HeatConfig(nx=nx, ny=ny, steps=41, k=8, px=px, py=py, arith="strict-fma", chunk_rows=32,
                     conv=True, interval=20, eps=0.0)
    s = Heat2DSolver(cfg, DEV).load(u0)
    s.run()
    assert np.array_equal(s.gather().numpy(), oracle(u0, 41, "fixed", "strict-fma"))


def test_strict_overflow_to_infinity_is_reproduced():
    # a neighbour sum that overflows float (3e38 + 3e38) is +inf in the reference arithmetic
    nx, ny = 64, 256
    u0 = np.full((nx, ny), np.float32(1.0), dtype=np.float32)
    u0[30:34, 100:104] = np.float32(3.0e38)
    s = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=16, k=8, arith="strict-fma"), DEV).load(u0)
    s.run()
    got, want = s.gather().numpy(), oracle(u0, 16, "fixed", "strict-fma")
    assert (~np.isfinite(want)).any()
    assert same_bits_or_nan(got, want)


# ------------------------------------------------- 8 columns per lane (strict arithmetic, k <= 4)
@pytest.mark.parametrize("arith", ["strict", "strict-fma"])
@pytest.mark.parametrize("k", [1, 2, 3, 4])
@pytest.mark.parametrize("boundary", ["fixed", "ghost-ring"])
def test_v8_layout_every_k_bit_exact(arith, k, boundary):
    nx, ny = 150, 777      # several 248-column strips, a shifted last strip, odd sizes
    u0 = rand_grid(nx, ny, k)
    u0[:, 300:320] = 0.0
    cfg = HeatConfig(nx=nx, ny=ny, steps=3 * k + 1, k=k, halo=8, boundary=boundary, arith=arith, chunk_rows=32,
                     layout="v8")
    s = Heat2DSolver(cfg, DEV).load(u0)
    assert s.engines[0].plan.layout == "v8"
    s.run()
    assert np.array_equal(s.gather().numpy(), oracle(u0, 3 * k + 1, boundary, arith))


@pytest.mark.parametrize("shape", [(10, 10), (80, 64), (640, 512), (1030, 2050), (97, 240), (64, 249), (33, 1000)])
def test_v8_layout_on_reference_sizes(shape):
    nx, ny = shape
    u0 = K.cpu_initial_grid(nx, ny, "exact")
    s = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=50, k=4, arith="strict-fma", layout="v8"), DEV).initialize()
    s.run()
    assert np.array_equal(s.gather().numpy(), oracle(u0, 50, "fixed", "strict-fma"))


@pytest.mark.parametrize("px,py", [(2, 2), (3, 1), (1, 3), (2, 4)])
@pytest.mark.parametrize("boundary", ["fixed", "ghost-ring"])
def test_v8_layout_virtual_tiles(px, py, boundary):
    # halo forwarding (row bands, packed column ghosts, corners) behind the 8-column task body
    for nx, ny in ((203, 517), (256, 2048)):
        u0 = rand_grid(nx, ny, 9)
        cfg = HeatConfig(nx=nx, ny=ny, steps=22, k=4, px=px, py=py, boundary=boundary, chunk_rows=16,
                         arith="strict-fma", layout="v8")
        s = Heat2DSolver(cfg, DEV).load(u0)
        assert s.engines[0].plan.layout == "v8"
        res = s.run()
        assert res.steps_done == 22
        assert np.array_equal(s.gather().numpy(), oracle(u0, 22, boundary, "strict-fma")), (nx, ny)


@pytest.mark.parametrize("px,py", [(1, 1), (2, 2)])
def test_v8_layout_convergence_and_fallback(px, py):
    nx, ny = 48, 600
    cfg = HeatConfig(nx=nx, ny=ny, steps=100000, k=4, px=px, py=py, boundary="ghost-ring", conv=True,
                     interval=20, eps=0.1, arith="strict-fma", layout="v8")
    s = Heat2DSolver(cfg, DEV).initialize()
    res = s.run()
    ref = K.cpu_initial_grid(nx, ny, "exact")
    done, _ = K.cpu_run(ref, 100000, N.make_phys(0.1, 0.1, "ghost-ring", "strict-fma"), "l2sq", 20, 0.1)
    assert res.converged and res.iterations == done
    assert np.array_equal(s.gather().numpy(), ref)
    # data outside the fast evaluation's domain: tasks are redone with the plain one
    u0 = rand_grid(200, 900, 22) - np.float32(5e3)
    u0[17, 33] = np.float32("inf")
    s = Heat2DSolver(HeatConfig(nx=200, ny=900, steps=20, k=4, px=px, py=py, arith="strict-fma", chunk_rows=32,
                                layout="v8"), DEV).load(u0)
    s.run()
    assert same_bits_or_nan(s.gather().numpy(), oracle(u0, 20, "fixed", "strict-fma"))


def test_v8_layout_pipelined_solve_and_fallback_to_v4():
    nx, ny = 1030, 2050
    u0 = rand_grid(nx, ny, 4)
    hin = torch.from_numpy(u0).pin_memory()
    hout = torch.empty((nx, ny), dtype=torch.float32).pin_memory()
    s = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=22, k=4, arith="strict-fma", layout="v8"), DEV)
    s.solve_pipelined(hin, hout, 22)
    assert np.array_equal(hout.numpy(), oracle(u0, 22, "fixed", "strict-fma"))
    # deeper sweeps and fp32 have no 8-column kernels: the plan falls back
    s = Heat2DSolver(HeatConfig(nx=64, ny=300, steps=8, k=8, arith="strict-fma", layout="v8"), DEV).initialize()
    assert s.engines[0].plan.layout == "v4"
    s = Heat2DSolver(HeatConfig(nx=64, ny=300, steps=8, k=4, arith="fp32", layout="v8"), DEV).initialize()
    assert s.engines[0].plan.layout in ("v4", "ab2", "ab4")


# ------------------------------------------------------------------------ failure paths
def test_linf_norm_propagates_nan():
    nx, ny = 64, 128
    a = torch.zeros((nx, ny), dtype=torch.float32, device=DEV)
    b = torch.zeros((nx, ny), dtype=torch.float32, device=DEV)
    a[13, 77] = float("nan")
    t = N.make_tile(nx, ny)
    assert torch.isnan(K.norm_diff(a, b, t, "linf")).item()
    assert torch.isnan(K.norm_diff(a, b, t, "l2sq")).item()
    # and inside the fused kernel: a field with a NaN never counts as converged
    u0 = rand_grid(nx, ny, 1)
    u0[20, 20] = np.float32("nan")
    cfg = HeatConfig(nx=nx, ny=ny, steps=41, k=8, conv=True, interval=20, eps=1e30, norm="linf")
    res = Heat2DSolver(cfg, DEV).load(u0).run()
    assert not res.converged and res.steps_done == 41 and np.isnan(res.last_norm)


def test_spin_watchdog_turns_a_silent_peer_into_an_error():
    """The GPU counterpart of MPI_Abort (grad1612_mpi_heat.c:55-63): a neighbour whose epoch
    flag never advances must not hang the kernel.  Tile 0 of a 2 x 1 virtual grid is launched
    for epoch 2 although its neighbour never ran epoch 1."""
    cfg = HeatConfig(nx=128, ny=256, steps=8, k=8, px=2, py=1, spin_timeout_s=0.01)
    s = Heat2DSolver(cfg, DEV).initialize()
    e0 = s.engines[0]
    e0.launch(0, 8, 2, "none")
    torch.cuda.synchronize()
    assert e0.fault() == 1 + N.DIRS.index("S")
    with pytest.raises(RuntimeError, match="direction S"):
        s._check_faults()
    # launches queued after the fault drain at once instead of spinning again
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    for ep in range(3, 40):
        e0.launch(ep & 1, 8, ep, "none")
    t1.record()
    t1.synchronize()
    assert t0.elapsed_time(t1) < 200.0     # 37 launches x 10 ms would be 370 ms
