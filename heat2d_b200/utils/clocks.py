"""Sample SM clocks and throttle reasons with nvidia-smi while a timed region runs.

The bench contract wants ``{"sm_mhz": median under load, "sm_max_mhz", "reasons"}``
taken DURING the timed region (B200_PROFILING.md, clocks line).
"""
from __future__ import annotations

import shutil
import statistics
import subprocess
import threading
from typing import List, Optional

QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
REASONS = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")


class ClockSampler:
    def __init__(self, gpu_index: int = 0, period_s: float = 0.1):
        self.gpu_index = gpu_index
        self.period_s = period_s
        self.samples: List[dict] = []
        self._stop = threading.Event()
        self._thr: Optional[threading.Thread] = None
        self.available = shutil.which("nvidia-smi") is not None

    def _once(self):
        try:
            out = subprocess.run(
                ["nvidia-smi", f"--query-gpu={QUERY}", "--format=csv,noheader,nounits", "-i",
                 str(self.gpu_index)], capture_output=True, text=True, timeout=5).stdout.strip()
            f = [x.strip() for x in out.split(",")]
            if len(f) < 8:
                return
            self.samples.append({
                "sm": float(f[1]), "max": float(f[2]),
                "power": float(f[3]) if f[3].replace(".", "", 1).isdigit() else float("nan"),
                "reasons": [r for r, v in zip(REASONS, f[4:8]) if v.lower().startswith("active")],
            })
        except Exception:
            pass

    def _loop(self):
        while not self._stop.is_set():
            self._once()
            self._stop.wait(self.period_s)

    def __enter__(self):
        if self.available:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thr is not None:
            self._thr.join(timeout=10)
        return False

    def summary(self) -> dict:
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = [s["sm"] for s in self.samples]
        # "under load": drop the idle-clock samples at the edges
        hi = max(sm)
        load = [x for x in sm if x >= 0.5 * hi] or sm
        reasons = sorted({r for s in self.samples for r in s["reasons"]})
        return {"sm_mhz": statistics.median(load), "sm_max_mhz": max(s["max"] for s in self.samples),
                "reasons": reasons, "samples": len(self.samples),
                "power_w_max": max((s["power"] for s in self.samples if s["power"] == s["power"]), default=None)}
