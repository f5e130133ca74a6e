"""The plan's task-sizing heuristic (host-only code of libheat2d_cuda.so; runs without a GPU,
where the SM count falls back to 148)."""
import ctypes as C

import pytest

from heat2d_b200 import _native as N


def chunk_rows(rows, cols, k=8, peers=False, layout=1):
    """layout 1 = v4 (128-column strips, 12 / 16 resident warps): the sweeps these numbers were
    fitted to; layout 3 = ab4 (256-column strips, 8 warps)."""
    lib = N.cuda()
    t = N.make_tile(rows, cols, 0, 0, rows, cols, 8, 32, (cols + 64 + 31) // 32 * 32)
    return lib.h2d_plan_chunk_rows_ex(C.byref(t), k, int(peers), 0, layout)


def waves(rows, cols, R, k=8):
    wv = 128 - 2 * ((k + 3) // 4 * 4)
    ns = -(-cols // wv)
    return -(-rows // R) * ns / (148 * (16 if k <= 5 else 12))


@pytest.mark.parametrize("rows,cols,lo,hi", [(32768, 32768, 450, 1024), (16384, 16384, 256, 512),
                                              (8192, 32768, 300, 400), (4096, 32768, 200, 260)])
def test_large_tiles_get_the_measured_best_chunks(rows, cols, lo, hi):
    # fitted to sweeps on B200 (profiles/ROOFLINE.md): within 1 % of the best chunk size
    R = chunk_rows(rows, cols)
    assert lo <= R <= hi, R
    w = waves(rows, cols, R)
    assert w <= round(w) + 1e-9 or w - int(w) > 0.9   # just under a whole number of waves


def test_small_grids_use_short_chunks():
    assert chunk_rows(640, 512) == 8        # one wave anyway: shortest tasks win (latency regime)
    assert chunk_rows(10, 10) == 10         # a single chunk
    assert chunk_rows(2560, 2048) <= 32


def test_plans_with_peers_keep_edge_tasks_short():
    # >= 2 waves per launch: the edge tasks (first wave) gate the neighbours' next launch.
    # (4 was the round-1 rule; on 8 B200 2 is 2-3 % faster: 30979 vs 30360 GLUP/s at 8 x 1,
    # profiles/scaling_sweep_n8.jsonl)
    for rows, cols in [(32768, 8192), (4096, 32768), (8192, 16384)]:
        R = chunk_rows(rows, cols, peers=True)
        assert waves(rows, cols, R) > 1.5
        assert R <= chunk_rows(rows, cols, peers=False)


def test_wide_strip_layout_fills_whole_waves_too():
    # ab4: 256-column strips (240 stored at k = 8), 8 resident warps per SM
    R = chunk_rows(32768, 32768, layout=3)
    ns = -(-32768 // 240)
    w = -(-32768 // R) * ns / (148 * 8)
    assert 200 <= R <= 2048 and (w <= round(w) + 1e-9 or w - int(w) > 0.85), (R, w)
    assert N.cuda().h2d_plan_strip_width(0, 3) == 256 and N.cuda().h2d_plan_strip_width(0, 1) == 128
    assert N.cuda().h2d_plan_strip_width(2, 3) == 128     # strict arithmetic always runs v4


def test_eight_column_layout_of_the_strict_modes():
    # v8: 256-column strips (248 stored), 12 resident warps; strict arithmetic, plans with kmax <= 4
    lib = N.cuda()
    assert lib.h2d_plan_strip_width(2, 4) == 256 and lib.h2d_plan_strip_width(1, 4) == 256
    assert lib.h2d_plan_strip_width(0, 4) == 128            # fp32 has no 8-column kernels: the default layout
    t = N.make_tile(32768, 32768, 0, 0, 32768, 32768, 8, 32, (32768 + 64 + 31) // 32 * 32)
    R = lib.h2d_plan_chunk_rows_ex(C.byref(t), 4, 0, 2, 4)
    ns = -(-32768 // 248)
    w = -(-32768 // R) * ns / (148 * 12)
    assert 200 <= R <= 2048 and (w <= round(w) + 1e-9 or w - int(w) > 0.85), (R, w)
    # a deeper plan cannot use it: the sizing falls back to 128-column strips
    assert lib.h2d_plan_chunk_rows_ex(C.byref(t), 8, 0, 2, 4) == lib.h2d_plan_chunk_rows_ex(C.byref(t), 8, 0, 2, 1)
