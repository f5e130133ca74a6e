"""Per-phase timers for the drivers (CLI, bench): wall clock on the host and, when a
device is given, CUDA events on the current stream.  The reference times exactly one
window (barrier .. end of the step loop, ``grad1612_mpi_heat.c:206-207,277-280``); the
phases here show what surrounds it (init, dumps, gather)."""
from __future__ import annotations

import time
from contextlib import contextmanager
from typing import Dict, Optional


class PhaseTimer:
    def __init__(self, device=None):
        self.device = device
        self.wall: Dict[str, float] = {}
        self.dev: Dict[str, float] = {}

    @contextmanager
    def phase(self, name: str):
        ev0 = ev1 = None
        if self.device is not None:
            import torch
            if torch.cuda.is_available():
                ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                ev0.record(torch.cuda.current_stream(self.device))
        t0 = time.perf_counter()
        try:
            yield self
        finally:
            if ev1 is not None:
                ev1.record(torch.cuda.current_stream(self.device))
                ev1.synchronize()
                self.dev[name] = self.dev.get(name, 0.0) + ev0.elapsed_time(ev1) / 1e3
            self.wall[name] = self.wall.get(name, 0.0) + time.perf_counter() - t0

    def as_dict(self) -> dict:
        return {"wall_s": dict(self.wall), "device_s": dict(self.dev)}
