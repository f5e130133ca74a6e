"""Per-phase timers for the drivers (CLI, bench): wall clock on the host and, when a
device is given, CUDA events on the current stream.  The reference times exactly one
window (barrier .. end of the step loop, ``grad1612_mpi_heat.c:206-207,277-280``); the
phases here show what surrounds it (init, dumps, gather).

With ``H2D_NVTX=1`` in the environment every phase (and every ``nvtx_range``) is also an
NVTX range, so a timeline tool shows the same names.  The reference's tracing was
link-time mpiP interposition (``Report.pdf`` p34-37); it has no in-source hooks."""
from __future__ import annotations

import os
import time
from contextlib import contextmanager
from typing import Dict


def nvtx_enabled() -> bool:
    return os.environ.get("H2D_NVTX", "0") not in ("", "0")


@contextmanager
def nvtx_range(name: str):
    """NVTX range around a block when H2D_NVTX=1 and CUDA is usable; otherwise free."""
    pushed = False
    if nvtx_enabled():
        try:
            import torch
            if torch.cuda.is_available():
                torch.cuda.nvtx.range_push(name)
                pushed = True
        except Exception:  # NVTX is a convenience, never a reason to fail a run
            pushed = False
    try:
        yield
    finally:
        if pushed:
            import torch
            torch.cuda.nvtx.range_pop()


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
            with nvtx_range(name):
                yield self
        finally:
            if ev1 is not None:
                ev1.record(torch.cuda.current_stream(self.device))
                ev1.synchronize()
                self.dev[name] = self.dev.get(name, 0.0) + ev0.elapsed_time(ev1) / 1e3
            self.wall[name] = self.wall.get(name, 0.0) + time.perf_counter() - t0

    def as_dict(self) -> dict:
        return {"wall_s": dict(self.wall), "device_s": dict(self.dev)}
