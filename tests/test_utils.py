"""Host utilities: clock sampler (fake nvidia-smi on PATH), NUMA cpulist parsing, phase timer."""
import os
import stat
import time

from heat2d_b200.utils.clocks import ClockSampler
from heat2d_b200.utils.numa import _parse_cpulist, bind_to_gpu_numa
from heat2d_b200.utils.timing import PhaseTimer, nvtx_range


def _fake_smi(tmp_path, lines):
    """An nvidia-smi that answers the sampler's query with the given csv lines in turn."""
    data = tmp_path / "answers.txt"
    data.write_text("\n".join(lines) + "\n")
    counter = tmp_path / "n"
    counter.write_text("0")
    exe = tmp_path / "nvidia-smi"
    exe.write_text(f"""#!/bin/sh
n=$(cat {counter}); total=$(wc -l < {data})
i=$(( n % total + 1 )); echo $(( n + 1 )) > {counter}
sed -n "${{i}}p" {data}
""")
    exe.chmod(exe.stat().st_mode | stat.S_IEXEC)
    return exe


def test_clock_sampler_reports_median_under_load_and_reasons(tmp_path, monkeypatch):
    _fake_smi(tmp_path, [
        "0, 210, 1965, 90.1, Not Active, Not Active, Not Active, Not Active",      # idle edge sample
        "0, 1950, 1965, 700.5, Not Active, Not Active, Not Active, Active",
        "0, 1920, 1965, 905.0, Not Active, Not Active, Not Active, Active",
        "0, 1965, 1965, [N/A], Not Active, Not Active, Not Active, Not Active",
    ])
    monkeypatch.setenv("PATH", f"{tmp_path}{os.pathsep}{os.environ['PATH']}")
    cs = ClockSampler(0, period_s=0.01)
    assert cs.available
    for _ in range(4):
        cs._once()
    s = cs.summary()
    assert s["samples"] == 4
    assert s["sm_mhz"] == 1950.0            # median of the three samples under load
    assert s["sm_max_mhz"] == 1965.0
    assert s["reasons"] == ["sw_power_cap"]
    assert s["power_w_max"] == 905.0


def test_clock_sampler_thread_and_garbage_output(tmp_path, monkeypatch):
    _fake_smi(tmp_path, ["garbage", "0, 1800, 1965, 500, Active, Not Active, Not Active, Not Active"])
    monkeypatch.setenv("PATH", f"{tmp_path}{os.pathsep}{os.environ['PATH']}")
    with ClockSampler(0, period_s=0.01) as cs:
        time.sleep(0.3)
    s = cs.summary()
    assert s["samples"] >= 1 and s["reasons"] == ["hw_slowdown"]


def test_clock_sampler_without_nvidia_smi(monkeypatch, tmp_path):
    monkeypatch.setenv("PATH", str(tmp_path))
    with ClockSampler(0) as cs:
        pass
    assert not cs.available
    assert cs.summary() == {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}


def test_cpulist_parsing_and_binding_never_raises():
    assert _parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert _parse_cpulist("") == set()
    info = bind_to_gpu_numa(0)            # no GPU here: reports the error, binds nothing
    assert info["device"] == 0 and info["bound"] is False


def test_phase_timer_accumulates_and_nvtx_is_free_without_cuda(monkeypatch):
    monkeypatch.setenv("H2D_NVTX", "1")
    t = PhaseTimer()
    for _ in range(2):
        with t.phase("a"):
            time.sleep(0.01)
    with t.phase("b"), nvtx_range("inner"):
        pass
    d = t.as_dict()
    assert d["wall_s"]["a"] >= 0.02 and "b" in d["wall_s"] and d["device_s"] == {}


def test_lazy_module_imports_on_first_use():
    import sys
    from heat2d_b200._lazy import LazyModule
    sys.modules.pop("colorsys", None)
    m = LazyModule("colorsys")
    assert "not loaded" in repr(m) and "colorsys" not in sys.modules
    assert m.rgb_to_hsv(1.0, 0.0, 0.0)[0] == 0.0
    assert "colorsys" in sys.modules and "(loaded)" in repr(m)


def test_build_uses_shipped_libraries_when_sources_are_absent(tmp_path):
    """An installed copy has the .so files but no csrc/: they must be used as they are."""
    from heat2d_b200 import build as B
    lib = tmp_path / "lib.so"
    lib.write_bytes(b"x")
    assert B._newer(lib, [tmp_path / "missing.cpp"])          # no sources -> up to date
    src = tmp_path / "a.cpp"
    src.write_text("int x;")
    os.utime(lib, (1, 1))
    assert not B._newer(lib, [src])                            # a newer source -> rebuild
    assert not B._newer(tmp_path / "absent.so", [src])
