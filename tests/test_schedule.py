import pytest

from heat2d_b200.config import ConfigError, HeatConfig, make_schedule


def test_plain_schedule():
    s = make_schedule(1000, 8)
    assert sum(l.k for l in s) == 1000 and len(s) == 125 and not any(l.check for l in s)
    # never a trailing shallow launch: equal depths, as few launches as k allows
    assert [l.k for l in make_schedule(10, 8)] == [5, 5]
    assert [l.k for l in make_schedule(20, 8)] == [7, 7, 6]
    assert [l.k for l in make_schedule(5, 8)] == [5]
    assert make_schedule(0, 8) == []


def test_conv_schedule_is_the_same_for_both_modes():
    a = make_schedule(100, 8, conv=True, interval=20, conv_mode="exact")
    b = make_schedule(100, 8, conv=True, interval=20, conv_mode="fused-k")
    assert a == b
    pos = 0
    for l in a:
        assert l.first_step == pos
        pos += l.k


def test_fused_conv_schedule_ends_launch_on_checked_step():
    s = make_schedule(100, 8, conv=True, interval=20, conv_mode="fused-k")
    assert sum(l.k for l in s) == 100
    ends = [l.first_step + l.k - 1 for l in s if l.check]
    assert ends == [0, 20, 40, 60, 80]
    assert all(1 <= l.k <= 8 for l in s)


def test_schedule_with_start_step():
    s = make_schedule(50, 8, conv=True, interval=20, conv_mode="exact", start_step=30)
    assert [l.first_step + l.k - 1 for l in s if l.check] == [40, 60]
    assert sum(l.k for l in s) == 50


def test_config_validation_messages():
    HeatConfig().validate()
    with pytest.raises(ConfigError, match="number of tasks must be equal to 4"):
        HeatConfig(px=2, py=2).validate(world_size=3)
    with pytest.raises(ConfigError, match="is not an integer"):
        HeatConfig(nx=10, ny=10, px=3, uneven=False).validate()
    with pytest.raises(ConfigError):
        HeatConfig(k=9).validate()
    with pytest.raises(ConfigError):
        HeatConfig(arith="fp16").validate()
    c = HeatConfig(cx_float=True)
    assert c.cx_eff != 0.1 and abs(c.cx_eff - 0.1) < 1e-8


def test_best_k_follows_the_measurements():
    from heat2d_b200 import best_k
    assert best_k("fp32", 1 << 30) == 8 and best_k("fp32", 1 << 20) == 8
    assert best_k("strict-fma", 32768 * 32768) == 4 and best_k("strict", 16384 * 16384) == 4
    assert best_k("strict-fma", 8192 * 16384) == 4      # an eighth of 32768^2: one tile of an 8-GPU run
    assert best_k("strict-fma", 8192 * 8192) == 8
