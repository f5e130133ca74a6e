"""Every sm_100a kernel against the CPU oracle (bit-exact: the oracle uses the
same operation order, SURVEY 4 "Kernel")."""
import os

import numpy as np
import pytest
import torch

from heat2d_b200 import _native as N
from heat2d_b200.config import HeatConfig
from heat2d_b200.models.heat import Heat2DSolver, TileField
from heat2d_b200.ops import kernels as K
from heat2d_b200.parallel.decomp import Decomposition

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def oracle(nx, ny, steps, boundary="fixed", arith="fp32", init="exact", u0=None):
    u = K.cpu_initial_grid(nx, ny, init) if u0 is None else u0.copy()
    K.cpu_run(u, steps, N.make_phys(0.1, 0.1, boundary, arith))
    return u


def rand_grid(nx, ny, seed=0):
    return (np.random.default_rng(seed).random((nx, ny), dtype=np.float32) * 1e4).astype(np.float32)


def test_library_loaded(native_cuda):
    assert b"heat2d" in native_cuda.h2d_version()
    assert "B200" in K.device_info(0) or "NVIDIA" in K.device_info(0)


@pytest.mark.parametrize("mode", ["exact", "ref-int32", "ref-uint32", "zero"])
@pytest.mark.parametrize("shape", [(10, 10), (640, 640), (700, 1000)])
def test_init_kernel(mode, shape):
    nx, ny = shape
    d = Decomposition(nx, ny, 1, 1)
    f = TileField(d.tile(0), nx, ny, 8, DEV)
    f.buf[0].fill_(-3.0)
    K.init_field(f.buf[0], f.tile, mode, True)
    want = np.zeros(tuple(f.buf[0].shape), dtype=np.float32)
    K.cpu_init(want, f.tile, mode, True)
    got = f.buf[0].cpu().numpy()
    t = f.tile
    w = t.cols + 2 * t.halo_c
    assert np.array_equal(got[:, :w], want[:, :w])


def test_init_kernel_tile_margins_hold_neighbour_values():
    nx, ny = 96, 160
    d = Decomposition(nx, ny, 2, 2)
    full = K.cpu_initial_grid(nx, ny, "exact")
    for spec in d.tiles:
        f = TileField(spec, nx, ny, 8, DEV)
        K.init_field(f.buf[0], f.tile, "exact", True)
        gx0, gx1, gy0, gy1, r0, c0 = f.padded_window()
        got = f.buf[0].cpu().numpy()[r0:r0 + gx1 - gx0, c0:c0 + gy1 - gy0]
        assert np.array_equal(got, full[gx0:gx1, gy0:gy1])


@pytest.mark.parametrize("arith", ["fp32", "strict", "strict-fma"])
@pytest.mark.parametrize("boundary", ["fixed", "ghost-ring"])
@pytest.mark.parametrize("shape", [(10, 10), (80, 64), (257, 391)])
def test_plain_kernel_bit_exact(arith, boundary, shape):
    nx, ny = shape
    u0 = rand_grid(nx, ny, 5)
    cfg = HeatConfig(nx=nx, ny=ny, steps=7, backend="plain", arith=arith, boundary=boundary, k=1)
    s = Heat2DSolver(cfg, DEV).load(u0)
    res = s.run()
    assert res.steps_done == 7
    assert np.array_equal(s.gather().numpy(), oracle(nx, ny, 7, boundary, arith, u0=u0))


@pytest.mark.parametrize("k", [1, 2, 3, 4, 5, 6, 7, 8])
@pytest.mark.parametrize("boundary", ["fixed", "ghost-ring"])
def test_tk_kernel_every_k_bit_exact(k, boundary):
    nx, ny = 150, 333      # several strips, partial last strip, odd sizes
    u0 = rand_grid(nx, ny, k)
    cfg = HeatConfig(nx=nx, ny=ny, steps=2 * k + 1, k=k, boundary=boundary, chunk_rows=32)
    s = Heat2DSolver(cfg, DEV).load(u0)
    res = s.run()
    assert res.steps_done == 2 * k + 1
    assert np.array_equal(s.gather().numpy(), oracle(nx, ny, 2 * k + 1, boundary, "fp32", u0=u0))


@pytest.mark.parametrize("arith", ["strict", "strict-fma"])
def test_tk_kernel_strict_modes(arith):
    nx, ny = 96, 200
    u0 = rand_grid(nx, ny, 11)
    cfg = HeatConfig(nx=nx, ny=ny, steps=24, k=8, arith=arith, chunk_rows=32)
    s = Heat2DSolver(cfg, DEV).load(u0)
    s.run()
    assert np.array_equal(s.gather().numpy(), oracle(nx, ny, 24, "fixed", arith, u0=u0))


@pytest.mark.parametrize("shape", [(10, 10), (80, 64), (640, 512), (1030, 2050)])
def test_tk_kernel_reference_sizes(shape):
    nx, ny = shape
    cfg = HeatConfig(nx=nx, ny=ny, steps=100, k=8, init="exact")
    s = Heat2DSolver(cfg, DEV).initialize()
    s.run()
    assert np.array_equal(s.gather().numpy(), oracle(nx, ny, 100))


def test_temporal_blocking_equals_single_steps():
    nx, ny = 300, 500
    u0 = rand_grid(nx, ny, 2)
    outs = []
    for k in (1, 4, 8):
        s = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=40, k=k), DEV).load(u0)
        s.run()
        outs.append(s.gather().numpy())
    assert np.array_equal(outs[0], outs[1]) and np.array_equal(outs[0], outs[2])


@pytest.mark.parametrize("px,py", [(2, 2), (2, 4), (3, 1), (1, 3), (4, 2)])
@pytest.mark.parametrize("boundary", ["fixed", "ghost-ring"])
def test_virtual_tiles_equal_single_tile(px, py, boundary):
    # P x Q result == 1 x 1 result bit for bit: the peer-store / flag
    # code path of the multi-GPU kernel, pointed at local buffers.
    nx, ny = 203, 317      # uneven splits
    u0 = rand_grid(nx, ny, 9)
    cfg = HeatConfig(nx=nx, ny=ny, steps=41, k=8, px=px, py=py, boundary=boundary, chunk_rows=16)
    s = Heat2DSolver(cfg, DEV).load(u0)
    res = s.run()
    assert res.steps_done == 41
    assert np.array_equal(s.gather().numpy(), oracle(nx, ny, 41, boundary, "fp32", u0=u0))


def test_virtual_tiles_init_needs_no_exchange():
    cfg = HeatConfig(nx=256, ny=384, steps=32, k=8, px=2, py=2)
    s = Heat2DSolver(cfg, DEV).initialize()
    s.run()
    assert np.array_equal(s.gather().numpy(), oracle(256, 384, 32))


@pytest.mark.parametrize("norm", ["l2sq", "linf"])
def test_norm_diff_kernel(norm):
    nx, ny = 123, 457
    d = Decomposition(nx, ny, 1, 1)
    f = TileField(d.tile(0), nx, ny, 8, DEV)
    a, b = rand_grid(nx, ny, 1), rand_grid(nx, ny, 2)
    f.owned(0).copy_(torch.from_numpy(a))
    f.owned(1).copy_(torch.from_numpy(b))
    got = float(K.norm_diff(f.buf[0], f.buf[1], f.tile, norm).item())
    d64 = a.astype(np.float64) - b.astype(np.float64)
    want = float((d64 ** 2).sum()) if norm == "l2sq" else float(np.abs(d64).max())
    assert got == pytest.approx(want, rel=1e-5)
    # second call on the same scratch-free path gives the same bits (deterministic order)
    assert float(K.norm_diff(f.buf[0], f.buf[1], f.tile, norm).item()) == got


@pytest.mark.parametrize("conv_mode", ["exact", "fused-k"])
@pytest.mark.parametrize("px,py", [(1, 1), (2, 2)])
def test_convergence_stops_like_the_oracle(conv_mode, px, py):
    nx, ny = 24, 24
    u0 = K.cpu_initial_grid(nx, ny, "exact")
    phys = N.make_phys(0.1, 0.1, "ghost-ring", "fp32")
    ref = u0.copy()
    done, last = K.cpu_run(ref, 100000, phys, norm="l2sq", interval=20, eps=0.1)
    assert 0 < done < 100000
    cfg = HeatConfig(nx=nx, ny=ny, steps=100000, k=8, px=px, py=py, boundary="ghost-ring", conv=True,
                     interval=20, eps=0.1, conv_mode=conv_mode)
    s = Heat2DSolver(cfg, DEV).initialize()
    res = s.run()
    assert res.converged
    assert res.last_norm == pytest.approx(last, rel=1e-3)
    if conv_mode == "exact":
        assert res.iterations == done and res.steps_done == done
        assert np.array_equal(s.gather().numpy(), ref)
    else:
        assert res.steps_done == done + 1
        nxt = ref.copy()
        K.cpu_run(nxt, 1, phys)
        assert np.array_equal(s.gather().numpy(), nxt)


def test_fused_norm_matches_standalone():
    nx, ny = 500, 700
    u0 = rand_grid(nx, ny, 21)
    cfg = HeatConfig(nx=nx, ny=ny, steps=1, k=1, conv=True, interval=1, eps=0.0, norm="linf")
    s = Heat2DSolver(cfg, DEV).load(u0)
    res = s.run()
    one = oracle(nx, ny, 1, u0=u0)
    assert res.last_norm == pytest.approx(float(np.abs(one.astype(np.float64) - u0).max()), rel=1e-6)


@pytest.mark.parametrize("shape", [(10, 10), (80, 64), (320, 256), (640, 512)])
@pytest.mark.parametrize("boundary", ["fixed", "ghost-ring"])
def test_resident_cluster_kernel(shape, boundary):
    nx, ny = shape
    assert K.resident_fits(nx, ny) >= 1
    cfg = HeatConfig(nx=nx, ny=ny, steps=100, backend="resident", boundary=boundary)
    s = Heat2DSolver(cfg, DEV).initialize()
    res = s.run()
    assert res.launches == 1
    assert np.array_equal(s.gather().numpy(), oracle(nx, ny, 100, boundary))


def test_resident_rejects_big_grid():
    assert K.resident_fits(4096, 4096) == 0


def test_bad_arguments_raise():
    d = Decomposition(64, 64, 1, 1)
    f = TileField(d.tile(0), 64, 64, 8, DEV)
    with pytest.raises(N.NativeError):
        K.Plan(f.tile, N.make_phys(), f.buf[0].data_ptr(), f.buf[1].data_ptr(), kmax=9)
    p = K.Plan(f.tile, N.make_phys(), f.buf[0].data_ptr(), f.buf[1].data_ptr(), kmax=4)
    with pytest.raises(N.NativeError):
        p.launch(0, 5, 1)
    p.close()


@pytest.mark.parametrize("px,py", [(1, 2), (2, 2), (1, 3)])
def test_plain_kernel_reads_ghosts_of_all_four_neighbours(px, py):
    # the compute half of the NCCL-halo baseline: one step per tile from a field
    # whose margins hold the neighbours' cells (cols not a multiple of 128 or 4)
    nx, ny = 90, 1401
    u0 = rand_grid(nx, ny, 33)
    want = oracle(nx, ny, 1, "ghost-ring", "fp32", u0=u0)
    d = Decomposition(nx, ny, px, py)
    phys = N.make_phys(0.1, 0.1, "ghost-ring", "fp32")
    got = np.zeros_like(u0)
    g = torch.from_numpy(u0)
    for spec in d.tiles:
        f = TileField(spec, nx, ny, 8, DEV)
        f.load_from_global(g, 0)
        K.step_plain(f.buf[0], f.buf[1], f.tile, phys)
        got[spec.x0:spec.x0 + spec.rows, spec.y0:spec.y0 + spec.cols] = f.owned(1).cpu().numpy()
    assert np.array_equal(got, want)


@pytest.mark.parametrize("cx,cy", [(0.07, 0.13), (0.2, 0.05)])
def test_tk_kernel_general_coefficients(cx, cy):
    # cx = cy = 0.1 takes the immediate-operand kernels; everything else the register form
    nx, ny = 200, 300
    u0 = rand_grid(nx, ny, 17)
    cfg = HeatConfig(nx=nx, ny=ny, steps=24, k=8, cx=cx, cy=cy, chunk_rows=32)
    s = Heat2DSolver(cfg, DEV).load(u0)
    s.run()
    ref = u0.copy()
    K.cpu_run(ref, 24, N.make_phys(cx, cy, "fixed", "fp32"))
    assert np.array_equal(s.gather().numpy(), ref)


def test_cli_gpu(tmp_path):
    import json
    import subprocess
    import sys
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "heat2d_b200", "--nx", "640", "--ny", "512", "--steps", "100",
                        "-k", "8", "--px", "2", "--py", "2", "--dump", "binary", "--debug", "--out-dir", str(tmp_path)],
                       cwd=root, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "Problem size:640x512" in r.stdout and "Each process will take: 320x256" in r.stdout
    assert "Elapsed time:" in r.stdout
    got = np.fromfile(str(tmp_path / "final_binary.dat"), dtype=np.float32).reshape(640, 512)
    assert np.array_equal(got, oracle(640, 512, 100))
    rec = json.loads((tmp_path / "result.json").read_text())
    assert rec["launches"] > 0 and rec["steps_done"] == 100


def test_config1_on_the_gpu_matches_the_reference_final_dat(tmp_path):
    """BASELINE config 1 through the GPU path: strict arithmetic, float coefficients,
    int32-wrapping init, 3 row strips (virtual tiles) -> final.dat byte-identical to the
    unmodified mpi_heat2Dn.c run with `mpiexec -n 4` on the MPI shim (or, where the
    reference cannot be built, to the CPU oracle that test_reference_shim.py ties to it)."""
    import sys
    from heat2d_b200.models.presets import preset
    from heat2d_b200.utils import dump as D
    cfg = preset("mpi_heat2Dn", nx=640, ny=640, steps=100, px=3, k=8, backend="fused", dump="none")
    s = Heat2DSolver(cfg, DEV).initialize()
    s.run()
    D.write_prtdat(str(tmp_path / "gpu_final.dat"), s.gather())
    ref = K.cpu_initial_grid(640, 640, "ref-int32")
    cf = float(np.float32(0.1))
    K.cpu_run(ref, 100, N.make_phys(cf, cf, "fixed", "strict"))
    D.write_prtdat(str(tmp_path / "oracle_final.dat"), ref)
    assert (tmp_path / "gpu_final.dat").read_bytes() == (tmp_path / "oracle_final.dat").read_bytes()
    try:
        sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline"))
        import ref_build as rb
        exe = rb.build_mpi("mpi_heat2Dn", {"NXPROB": 640, "NYPROB": 640, "STEPS": 100}, "-O2")
        rb.run_mpi(exe, 4, tmp_path)
    except Exception as e:  # reference not staged / no gcc on this box
        pytest.skip(f"reference run unavailable: {e}")
    assert (tmp_path / "final.dat").read_bytes() == (tmp_path / "gpu_final.dat").read_bytes()


def test_strict_mode_handles_zero_subnormal_and_huge_values():
    # the strict kernels convert float -> double with integer instructions for normal
    # numbers and with the hardware conversion otherwise: exercise both, bit-exactly
    nx, ny = 64, 160
    rng = np.random.default_rng(5)
    u0 = rng.random((nx, ny), dtype=np.float32)
    u0[:, :40] *= np.float32(1e-42)            # subnormals
    u0[:, 40:80] = 0.0
    u0[:, 80:120] *= np.float32(3e37)
    u0[10:20, :] *= np.float32(-1.0)
    for arith in ("strict", "strict-fma"):
        s = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=16, k=8, arith=arith, boundary="ghost-ring",
                                    chunk_rows=16), DEV).load(u0)
        s.run()
        assert np.array_equal(s.gather().numpy(), oracle(nx, ny, 16, "ghost-ring", arith, u0=u0))


def test_halo_wait_statistics_are_exposed():
    s = Heat2DSolver(HeatConfig(nx=256, ny=384, steps=32, k=8, px=2, py=2), DEV).initialize()
    s.run()
    us, n = s.halo_wait_stats()
    # ranks emulated on one stream never block: every flag is already published
    assert us == 0 and n == 0


@pytest.mark.parametrize("shape,steps,k,block,skew", [((700, 900), 83, 8, 96, 32), ((1030, 2050), 41, 5, 0, 0),
                                                      ((900, 1300), 42, 4, 0, 0), ((515, 777), 20, 4, 64, 0),
                                                      ((400, 600), 31, 3, 0, 0),
                                                      ((257, 391), 24, 8, 64, 9), ((300, 500), 0, 8, 0, 0)])
def test_pipelined_upload_solve_download(shape, steps, k, block, skew):
    # blocks run all their launches before the next block starts (skewed windows):
    # the result must not depend on the block size or the skew
    nx, ny = shape
    u0 = rand_grid(nx, ny, 77)
    hin = torch.from_numpy(u0).pin_memory()
    hout = torch.empty((nx, ny), dtype=torch.float32).pin_memory()
    s = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=steps, k=k, chunk_rows=32), DEV)
    res = s.solve_pipelined(hin, hout, steps, block_rows=block, skew_rows=skew)
    assert res.steps_done == steps
    want = oracle(nx, ny, steps, u0=u0)
    assert np.array_equal(hout.numpy(), want)
    assert np.array_equal(s.gather().numpy(), want)     # the device copy agrees and the parity is right
    # and the solver can keep going from there with ordinary launches
    s.run(16)
    assert np.array_equal(s.gather().numpy(), oracle(nx, ny, steps + 16, u0=u0))


def test_solve_api_uses_the_pipelined_path_for_pinned_input():
    from heat2d_b200 import solve
    nx, ny = 400, 640
    u0 = rand_grid(nx, ny, 3)
    res, field = solve(HeatConfig(nx=nx, ny=ny, steps=30, k=8), u0=torch.from_numpy(u0).pin_memory())
    assert field.is_pinned() and res.steps_done == 30
    assert np.array_equal(field.numpy(), oracle(nx, ny, 30, u0=u0))
    res2, field2 = solve(HeatConfig(nx=nx, ny=ny, steps=30, k=8), u0=u0)      # ordinary path, numpy in
    assert np.array_equal(field2.numpy(), field.numpy())


@pytest.mark.parametrize("py,shape,steps,k,block", [(2, (1500, 512), 83, 8, 0), (3, (900, 768), 41, 5, 128),
                                                    (4, (2200, 1024), 24, 8, 256), (2, (1100, 640), 42, 4, 0),
                                                    (3, (800, 900), 20, 4, 96),
                                                    # 228 columns per tile: the last task strip owns 4 < k
                                                    # columns, the strip before it reads the E ghosts too
                                                    (2, (1200, 456), 50, 8, 0)])
def test_pipelined_solve_on_virtual_column_strips(py, shape, steps, k, block):
    """1 x P tiles in one process: upload / window launches with W-E halo exchange / download,
    then an ordinary run() from the state it left (flags, epochs, ghost margins)."""
    nx, ny = shape
    u0 = rand_grid(nx, ny, py)
    hin = torch.from_numpy(u0).pin_memory()
    hout = torch.empty((nx, ny), dtype=torch.float32).pin_memory()
    s = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=steps, k=k, px=1, py=py, chunk_rows=48, layout="v4"), DEV)
    res = s.solve_pipelined(hin, hout, block_rows=block)
    torch.cuda.synchronize()
    assert res.steps_done == steps
    assert np.array_equal(hout.numpy(), oracle(nx, ny, steps, u0=u0))
    s.run(2 * k + 1)
    assert np.array_equal(s.gather().numpy(), oracle(nx, ny, steps + 2 * k + 1, u0=u0))
