"""Cost models (utils/perfmodel.py): the report's cluster model and the B200 three-roof model."""
import math

import pytest

from heat2d_b200.utils.perfmodel import B200Model, ClusterModel, roofline_fraction


def test_cluster_model_blocks_beat_strips():
    # The argument of Report.pdf p9-14: same rank count, 2-D blocks move less halo.
    m = ClusterModel()
    strips = m.step_us(2560, 2048, 16, 1)
    blocks = m.step_us(2560, 2048, 4, 4)
    assert blocks < strips
    assert m.speedup(2560, 2048, 1, 1) == pytest.approx(1.0)
    # one rank: pure compute, t_c per point
    assert m.run_s(2560, 2048, 1, 1, 1000) == pytest.approx(0.025e-6 * 2560 * 2048 * 1000)


def test_cluster_model_speedup_saturates():
    m = ClusterModel()
    s = [m.speedup(2560, 2048, p, p) for p in (1, 2, 4, 8)]
    assert s == sorted(s)
    assert s[-1] < 64  # communication keeps it below linear


def test_b200_roofs_k1_is_hbm_bound():
    m = B200Model(hbm_gbs=6551.0)
    b = m.bound_glups(16384, 16384, 1, blocking_overhead=False)
    assert b["limiter"] == "hbm"
    assert b["glups"] == pytest.approx(6551.0 / 8.0, rel=1e-6)


def test_b200_roofs_k8_is_pipe_bound():
    m = B200Model(hbm_gbs=6551.0)
    algo = m.bound_glups(32768, 32768, 8, blocking_overhead=False)
    # 1 B of HBM traffic per update (6.55 TLUP/s) sits just under the pipe roof (6.98)
    assert algo["limiter"] == "hbm"
    assert algo["glups"] == pytest.approx(6551.0, rel=1e-6)
    assert algo["pipe_us"] == pytest.approx(32768 * 32768 * 8 * 5 / 34.9e12 * 1e6, rel=1e-9)
    real = m.bound_glups(32768, 32768, 8, chunk_rows=683)
    assert real["limiter"] == "pipe"
    assert 5600 < real["glups"] < 6200                                         # profiles/ROOFLINE.md: 5.9
    # smaller chunks refill the pipeline more often
    assert m.bound_glups(32768, 32768, 8, chunk_rows=150)["glups"] < real["glups"]


def test_b200_small_grid_is_launch_bound():
    b = B200Model().bound_glups(80, 64, 8)
    assert b["limiter"] == "launch"


def test_nvlink_roof_is_negligible_for_row_strips():
    c = B200Model().launch_costs_us(4096, 32768, 8, 150, row_neighbours=2)
    assert c["nvlink_us"] < 0.05 * max(c["pipe_us"], c["hbm_us"])


def test_roofline_fraction_whole_job():
    r = roofline_fraction(26139.0, 32768, 32768, 8, 1, 8, chunk_rows=150)
    assert 0.4 < r["fraction_with_blocking"] < 1.0
    assert r["fraction_algorithm"] < r["fraction_with_blocking"]
    assert r["k1_hbm_bound_glups"] == pytest.approx(B200Model.from_measured().hbm_gbs)
    assert math.isfinite(r["bound_glups_algorithm"])
