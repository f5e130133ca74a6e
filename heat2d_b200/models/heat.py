"""The heat-diffusion model: tile storage, per-tile engine, single-process solver.

``Heat2DSolver`` is the public entry point a user of the reference switches to:
it plays the role of ``main()`` in all four reference programs
(``grad1612_mpi_heat.c:27-316``, ``grad1612_cuda_heat.cu:64-94``,
``mpi_heat2Dn.c:46-218``): allocate, initialise, (dump), timed step loop with
optional convergence test, (dump).  In one process it drives either one tile
(``px = py = 1``) or P x Q *virtual tiles* on one GPU — the same kernels, flag
protocol and peer stores as the multi-GPU path, pointed at local buffers and
issued rank after rank on one stream.  The one-process-per-GPU version is
:class:`heat2d_b200.parallel.dist.DistributedHeat2DSolver`.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional

import numpy as np
import torch

from .. import _native as N
from ..config import HeatConfig, Launch, make_schedule
from ..ops import kernels as K
from ..parallel.decomp import Decomposition, OPPOSITE, TileSpec
from ..utils.timing import nvtx_range

COL_MARGIN = 32  # floats; keeps owned column 0 128-byte aligned and >= roundup4(KMAX)
COL_PACK = 8     # floats per row of a packed column-ghost buffer (TK_CW)


def _round_up(v: int, m: int) -> int:
    return (v + m - 1) // m * m


def padded_geometry(spec: TileSpec, nx: int, ny: int, halo: int) -> N.Tile:
    halo_c = max(COL_MARGIN, _round_up(halo, 4))
    pitch = _round_up(spec.cols + 2 * halo_c, 32)
    return N.Tile(nx, ny, spec.x0, spec.y0, spec.rows, spec.cols, halo, halo_c, pitch)


def _layout_code(cfg: HeatConfig) -> int:
    """h2d_plan_desc_t::layout for a configuration (the dataflow launch exists for v4 only)."""
    from ..config import LAYOUTS
    return LAYOUTS["v4"] if (cfg.dataflow and cfg.layout == "auto") else LAYOUTS[cfg.layout]


def padded_bytes(tile: N.Tile) -> int:
    return (tile.rows + 2 * tile.halo_r) * tile.pitch * 4


class RawCudaArray:
    """Expose a raw device address to torch through __cuda_array_interface__."""

    def __init__(self, ptr: int, shape, typestr: str, owner=None):
        self._owner = owner
        self.__cuda_array_interface__ = {
            "shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
            "version": 2, "strides": None,
        }


def tensor_from_ptr(ptr: int, shape, dtype: torch.dtype, device, owner=None) -> torch.Tensor:
    typestr = {torch.float32: "<f4", torch.int32: "<i4", torch.uint8: "|u1", torch.int64: "<i8"}[dtype]
    return torch.as_tensor(RawCudaArray(ptr, shape, typestr, owner), device=device)


class TileField:
    """Double-buffered padded storage of one tile.

    ``buf[p]`` has shape ``(rows + 2*halo_r, pitch)``; ``owned(p)`` is the strided
    view of the cells this tile owns.  Margins are ghost cells (neighbour data
    or zeros outside the grid) — the k-wide successor of the 1-cell ghost ring
    of ``grad1612_mpi_heat.c:47-52,150-161``.
    """

    def __init__(self, spec: TileSpec, nx: int, ny: int, halo: int, device,
                 storage: Optional[List[torch.Tensor]] = None, colbuf=None):
        self.spec = spec
        self.tile = padded_geometry(spec, nx, ny, halo)
        self.device = torch.device(device)
        shape = (spec.rows + 2 * halo, self.tile.pitch)
        if storage is None:
            self.buf = [torch.zeros(shape, dtype=torch.float32, device=self.device) for _ in range(2)]
        else:
            self.buf = storage
            for b in self.buf:
                assert tuple(b.shape) == shape and b.dtype == torch.float32 and b.is_contiguous()
        # packed column-ghost buffers [side W/E][parity], rows x 8 floats each (csrc/tk_task.cuh):
        # the W / E neighbours write their edge columns here as one contiguous stream
        self.colbuf = colbuf if colbuf is not None else [
            [torch.zeros((spec.rows, COL_PACK), dtype=torch.float32, device=self.device) for _ in range(2)]
            for _ in range(2)]

    def owned(self, parity: int) -> torch.Tensor:
        t = self.tile
        return self.buf[parity][t.halo_r:t.halo_r + t.rows, t.halo_c:t.halo_c + t.cols]

    def padded_window(self):
        """Global index ranges covered by the padded array, clipped to the grid:
        (gx0, gx1, gy0, gy1, local row offset, local col offset)."""
        t = self.tile
        gx0, gx1 = max(0, t.x0 - t.halo_r), min(t.nx, t.x0 + t.rows + t.halo_r)
        gy0, gy1 = max(0, t.y0 - t.halo_c), min(t.ny, t.y0 + t.cols + t.halo_c)
        return gx0, gx1, gy0, gy1, gx0 - (t.x0 - t.halo_r), gy0 - (t.y0 - t.halo_c)

    def load_from_global(self, grid: torch.Tensor, parity: int = 0):
        """Copy this tile's part of a full host/device grid, ghosts included."""
        gx0, gx1, gy0, gy1, r0, c0 = self.padded_window()
        self.buf[parity].zero_()
        src = grid[gx0:gx1, gy0:gy1]
        self.buf[parity][r0:r0 + (gx1 - gx0), c0:c0 + (gy1 - gy0)].copy_(src, non_blocking=True)


@dataclass
class RunResult:
    steps_done: int            # number of updates applied to the returned field
    iterations: int            # what the reference prints as "Exiting after %d iterations"
    converged: bool
    last_norm: float
    elapsed_s: float           # device time of the step loop (max over ranks when distributed)
    launches: int
    checks: int
    nx: int
    ny: int
    steps_run: int = -1        # updates applied by THIS call (steps_done also counts earlier calls / start_step)

    @property
    def glups(self) -> float:
        """Throughput of this call: grid points x updates applied by it / its device time."""
        if self.elapsed_s <= 0:
            return float("nan")
        n = self.steps_run if self.steps_run >= 0 else self.steps_done
        return self.nx * self.ny * max(n, 0) / self.elapsed_s / 1e9


class TileEngine:
    """One tile's plan plus the bookkeeping the flag protocol needs."""

    def __init__(self, field: TileField, phys: N.Phys, kmax: int, chunk_rows: int, world: int,
                 rank: int, flags_in: torch.Tensor, vals: torch.Tensor, nflags: torch.Tensor,
                 spin_timeout_s: float = 10.0, layout: int = 0, team: int = 0):
        self.field = field
        self.team = team
        self.phys = phys
        self.kmax = kmax
        self.chunk_rows = chunk_rows
        self.world, self.rank = world, rank
        self.flags_in, self.vals, self.nflags = flags_in, vals, nflags
        self.spin_timeout_s = spin_timeout_s
        self.layout = layout
        dev = field.device
        self.norm_total = torch.zeros(1, dtype=torch.float32, device=dev)
        # run-control word: 0 = iterating, q = check number q stopped the run
        self.ctl = torch.zeros(1, dtype=torch.int32, device=dev)
        self.plan: Optional[K.Plan] = None
        self.vals_ptrs = None
        self.flags_ptrs = None

    def build(self, peers: Dict[str, Optional[K.PeerRef]], all_vals_ptrs: List[int],
              all_nflags_ptrs: List[int]):
        f = self.field
        self.plan = K.Plan(f.tile, self.phys, f.buf[0].data_ptr(), f.buf[1].data_ptr(), self.kmax,
                           self.chunk_rows, peers=peers, flags_in_ptr=self.flags_in.data_ptr(),
                           spin_timeout_s=self.spin_timeout_s, device=f.device, layout=self.layout, team=self.team,
                           colbuf_in=[[f.colbuf[side][par].data_ptr() if peers.get(d) is not None else 0
                                       for par in (0, 1)] for side, d in ((0, "W"), (1, "E"))])
        self.vals_ptrs = torch.tensor(all_vals_ptrs, dtype=torch.int64, device=f.device)
        self.flags_ptrs = torch.tensor(all_nflags_ptrs, dtype=torch.int64, device=f.device)
        self._fault = tensor_from_ptr(self.plan.fault_ptr, (1,), torch.int32, f.device)
        self._stats = tensor_from_ptr(self.plan.stats_ptr, (2,), torch.int64, f.device)

    def launch(self, parity: int, k: int, epoch: int, norm: str, run_if: int = 0):
        self.plan.launch(parity, k, epoch, norm, self.ctl.data_ptr(), run_if)

    def finish(self, seq: int, norm: str, eps: float, phase: int):
        """phase 0 = publish, 1 = combine, 2 = publish + wait + combine.  On
        convergence the kernel stores `seq` into the run-control word."""
        lib = N.cuda()
        dev = self.field.device
        with torch.cuda.device(dev):
            N.check(lib, lib.h2d_cuda_norm_finish_ctl(
                self.plan.norm_ptr, self.vals_ptrs.data_ptr(), self.flags_ptrs.data_ptr(), self.rank,
                self.world, seq & 0xFFFFFFFF, N.NORM[norm], float(eps), self.norm_total.data_ptr(),
                self.ctl.data_ptr(), phase, int(self.spin_timeout_s * 1e9), self.plan.fault_ptr,
                int(torch.cuda.current_stream(dev).cuda_stream)))

    def device_barrier(self, seq: int):
        """Queue a one-block kernel that returns only after EVERY rank's stream has reached
        the same point (the all-to-all of the norm finish with a threshold nothing meets).
        It aligns the start of a timed region across GPUs to a few microseconds, where a host
        barrier leaves the ranks' launch skew inside the window (MPI_Barrier before
        MPI_Wtime, grad1612_mpi_heat.c:206-207)."""
        lib = N.cuda()
        dev = self.field.device
        if not hasattr(self, "_bar_stop"):
            self._bar_stop = torch.zeros(1, dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            N.check(lib, lib.h2d_cuda_norm_finish(
                self.plan.norm_ptr, self.vals_ptrs.data_ptr(), self.flags_ptrs.data_ptr(), self.rank,
                self.world, seq & 0xFFFFFFFF, N.NORM["linf"], float("-inf"), self.norm_total.data_ptr(),
                self._bar_stop.data_ptr(), 2, int(self.spin_timeout_s * 1e9), self.plan.fault_ptr,
                int(torch.cuda.current_stream(dev).cuda_stream)))

    def fault(self) -> int:
        return int(self._fault.item())

    def strict_fallbacks(self):
        """Tasks of the last three launches that left the fast strict evaluation (3 counters)."""
        t = tensor_from_ptr(self.plan.strict_fallback_ptr, (3,), torch.int32, self.field.device)
        return [int(x) for x in t.tolist()]

    def halo_wait_stats(self):
        """(microseconds edge tasks spent blocked on neighbour flags, number of blocking waits)
        accumulated since the plan was created."""
        ns, n = (int(x) for x in self._stats.tolist())
        return ns / 1e3, n


class Heat2DSolver:
    """Single-process solver: one GPU, one tile or P x Q virtual tiles.

    >>> s = Heat2DSolver(HeatConfig(nx=640, ny=512, steps=1000, k=8))
    >>> s.initialize(); res = s.run(); u = s.gather()
    """

    def __init__(self, cfg: HeatConfig, device="cuda:0"):
        self.cfg = cfg.validate()
        self.device = torch.device(device)
        if self.cfg.backend == "cpu":
            raise ValueError("use heat2d_b200.models.cpu.CpuSolver for the cpu backend")
        self.decomp = Decomposition(cfg.nx, cfg.ny, cfg.px, cfg.py, cfg.uneven)
        self.phys = N.make_phys(cfg.cx_eff, cfg.cy_eff, cfg.boundary, cfg.arith)
        self.world = self.decomp.world
        if self.world > 1 and cfg.k > self.decomp.min_extent():
            raise ValueError(f"k = {cfg.k} exceeds the smallest tile extent {self.decomp.min_extent()}")
        self.parity = 0
        self.epoch = 0
        self.norm_seq = 0
        self.steps_done = cfg.start_step
        self.fields: List[TileField] = []
        self.engines: List[TileEngine] = []
        self._resident_buf = None
        self._build()

    # ------------------------------------------------------------------ build
    def _build(self):
        cfg, dev = self.cfg, self.device
        for spec in self.decomp.tiles:
            self.fields.append(TileField(spec, cfg.nx, cfg.ny, cfg.halo, dev))
        if cfg.backend in ("plain", "resident"):
            if self.world != 1:
                raise ValueError(f"backend {cfg.backend!r} needs px = py = 1 in a single process")
            return
        w = self.world
        for r, f in enumerate(self.fields):
            flags_in = torch.zeros(64, dtype=torch.int32, device=dev)
            vals = torch.zeros(2 * w, dtype=torch.float32, device=dev)
            nflags = torch.zeros(2 * w, dtype=torch.int32, device=dev)
            self.engines.append(TileEngine(f, self.phys, max(cfg.k, 1), cfg.chunk_rows, w, r, flags_in,
                                           vals, nflags, cfg.spin_timeout_s, _layout_code(cfg), int(cfg.team == "on")))
        all_vals = [e.vals.data_ptr() for e in self.engines]
        all_nfl = [e.nflags.data_ptr() for e in self.engines]
        for r, e in enumerate(self.engines):
            peers: Dict[str, Optional[K.PeerRef]] = {}
            for d, nb in self.decomp.tile(r).neighbours.items():
                if nb is None:
                    peers[d] = None
                    continue
                nf = self.fields[nb]
                # I publish my epoch into the slot of the neighbour's flags_in that
                # is indexed by the direction I am in as seen from there.
                slot = N.DIRS.index(OPPOSITE[d])
                # my W band lands in the west neighbour's EAST packed buffer and vice versa
                side = {"W": 1, "E": 0}.get(d)
                cb = [nf.colbuf[side][par].data_ptr() for par in (0, 1)] if side is not None else [0, 0]
                peers[d] = K.PeerRef(nf.buf[0].data_ptr(), nf.buf[1].data_ptr(),
                                     self.engines[nb].flags_in.data_ptr() + 4 * slot,
                                     nf.tile.x0, nf.tile.y0, nf.tile.halo_r, nf.tile.halo_c, nf.tile.pitch,
                                     cb[0], cb[1])
            e.build(peers if w > 1 else {}, all_vals, all_nfl)

    # ------------------------------------------------------------- initialise
    def initialize(self, mode: Optional[str] = None):
        """Initial condition, evaluated tile-locally including the ghost margins
        (``grad1612_mpi_heat.c:163-168``: every rank fills its own tile, no scatter)."""
        mode = mode or self.cfg.init
        for f in self.fields:
            f.buf[1].zero_()
            K.init_field(f.buf[0], f.tile, mode, True)
        self.parity = 0
        self.steps_done = self.cfg.start_step
        self._seed_colbufs()
        return self

    def seed_colbufs(self):
        """Call after writing a padded array (`fields[i].buf[parity]`) directly."""
        self._seed_colbufs()

    def _seed_colbufs(self):
        """The padded arrays were just filled by something other than the neighbours' launches:
        bring the packed column-ghost buffers of the current parity in line with the margins."""
        for e in self.engines:
            if e.plan is not None:
                e.plan.seed_colbuf(self.parity)

    def load(self, grid, parity: int = 0):
        """Scatter a full nx x ny grid (host or device tensor / ndarray) to the tiles."""
        if isinstance(grid, np.ndarray):
            grid = torch.from_numpy(grid)
        assert tuple(grid.shape) == (self.cfg.nx, self.cfg.ny), "grid shape mismatch"
        for f in self.fields:
            f.buf[1 - parity].zero_()
            f.load_from_global(grid, parity)
        self.parity = parity
        self._seed_colbufs()
        return self

    # -------------------------------------------------------------------- run
    def schedule(self, steps: Optional[int] = None) -> List[Launch]:
        cfg = self.cfg
        steps = cfg.steps if steps is None else steps
        k = 1 if cfg.backend == "plain" else cfg.k
        return make_schedule(steps, k, cfg.conv, cfg.interval, cfg.conv_mode, self.steps_done)

    def _finish_all(self, seq: int, norm: str):
        """Cross-tile finish of check number `seq` (no host synchronisation)."""
        cfg = self.cfg
        if len(self.engines) == 1:
            self.engines[0].finish(seq, norm, cfg.eps, 2)
        else:  # ranks emulated as consecutive launches: publish all, then combine all
            for e in self.engines:
                e.finish(seq, norm, cfg.eps, 0)
            for e in self.engines:
                e.finish(seq, norm, cfg.eps, 1)

    def run(self, steps: Optional[int] = None) -> RunResult:
        cfg = self.cfg
        steps = cfg.steps if steps is None else steps
        with nvtx_range(f"heat2d.run[{cfg.backend} k={cfg.k} steps={steps}]"):
            if cfg.backend == "resident":
                return self._run_resident(steps)
            if cfg.backend == "plain":
                return self._run_plain(steps)
            return self._run_fused(steps)

    def _run_fused(self, steps: int) -> RunResult:
        """Queue the whole run without ever synchronising with the device.

        Convergence is decided on the device: the finish kernel of check q writes q
        into every rank's run-control word, after which every queued launch is a
        no-op.  ``conv_mode='exact'`` additionally queues, after each k-step check
        launch, a (k-1)-step launch from the same source that runs only if that
        check converged: it regenerates the pre-step field the reference returns
        (it breaks out of the loop before flipping buffers,
        ``grad1612_mpi_heat.c:269-273``).  Both share one epoch number: at most one
        of "conditional launch" and "next ordinary launch" ever executes.
        """
        cfg = self.cfg
        sched = self.schedule(steps)
        norm = cfg.norm if cfg.conv else "none"
        ev0 = torch.cuda.Event(enable_timing=True)
        ev1 = torch.cuda.Event(enable_timing=True)
        plans = [e.plan for e in self.engines]
        ctls = [e.ctl.data_ptr() for e in self.engines]
        records = {}   # seq -> (parity before, k, first_step, epoch of the check launch, launches so far)
        launches = 0
        parity, epoch, done = self.parity, self.epoch, self.steps_done
        done0 = done
        with torch.cuda.device(self.device):
            if cfg.conv:
                for e in self.engines:
                    e.ctl.zero_()
            ev0.record()
            i = 0
            while i < len(sched):
                L = sched[i]
                if not (L.check and cfg.conv):
                    j = i
                    while j < len(sched) and sched[j].k == L.k and not (sched[j].check and cfg.conv):
                        j += 1
                    n = j - i
                    if cfg.dataflow and len(plans) == 1 and self.world == 1 and not cfg.conv and n > 1:
                        plans[0].run_dataflow(parity, L.k, n)
                        launches += 1
                    else:
                        K.Plan.run_many(plans, parity, L.k, n, epoch + 1, ctls if cfg.conv else None)
                        launches += n * len(plans)
                    epoch += n
                    parity ^= n & 1
                    done += n * L.k
                    i = j
                    continue
                epoch += 1
                self.norm_seq += 1
                seq = self.norm_seq
                for e in self.engines:
                    e.launch(parity, L.k, epoch, norm, 0)
                launches += len(plans)
                self._finish_all(seq, norm)
                records[seq] = (parity, L.k, L.first_step, epoch, launches)
                if cfg.conv_mode == "exact" and L.k > 1:
                    for e in self.engines:
                        e.launch(parity, L.k - 1, epoch + 1, "none", seq)
                parity ^= 1
                done += L.k
                i += 1
            ev1.record()
            ev1.synchronize()
        elapsed = ev0.elapsed_time(ev1) / 1e3
        converged, last_norm, checks, iterations = False, float("nan"), len(records), None
        if cfg.conv and records:
            q = int(self.engines[0].ctl.item())
            last_norm = float(self.engines[0].norm_total.item())
            if q != 0:
                converged = True
                p0, k, first, ep, nl = records[q]
                checks = sum(1 for s in records if s <= q)
                launches = nl
                if cfg.conv_mode == "exact":
                    # field after `first + k - 1` updates; the reference reports that index
                    done = first + k - 1
                    parity = p0 if k == 1 else 1 - p0
                    epoch = ep if k == 1 else ep + 1
                    launches += 0 if k == 1 else len(plans)
                    iterations = done
                else:
                    done = first + k
                    parity = 1 - p0
                    epoch = ep
                    iterations = done
        self.parity, self.epoch, self.steps_done = parity, epoch, done
        self._check_faults()
        if iterations is None:
            iterations = self.steps_done
        return RunResult(self.steps_done, iterations, converged, last_norm, elapsed, launches, checks,
                         cfg.nx, cfg.ny, self.steps_done - done0)

    def solve_pipelined(self, host_in: torch.Tensor, host_out: torch.Tensor, steps: Optional[int] = None,
                        block_rows: int = 0, skew_rows: int = 0) -> RunResult:
        """End-to-end solve of a field that lives in (pinned) host memory: upload,
        `steps` steps, download, with the three phases overlapped block by block
        (one tile, no convergence test).  `elapsed_s` covers all three."""
        cfg = self.cfg
        steps = cfg.steps if steps is None else steps
        if cfg.conv:
            raise ValueError("solve_pipelined has no convergence test; use run()")
        if len(self.engines) > 1 and cfg.px == 1 and cfg.backend == "fused":
            # 1 x P virtual tiles: every window is issued for every tile in turn; host_in / host_out
            # are the whole fields and each tile reads its ghost columns straight from host_in
            plans = [e.plan for e in self.engines]
            l0 = self.kernel_launches
            ev0 = torch.cuda.Event(enable_timing=True)
            ev1 = torch.cuda.Event(enable_timing=True)
            with torch.cuda.device(self.device):
                ev0.record()
                self.parity, used = K.Plan.solve_pipelined_many(
                    plans, host_in, host_out, [f.tile.y0 for f in self.fields], steps, cfg.k, self.epoch + 1,
                    block_rows, skew_rows)
                ev1.record()
                ev1.synchronize()
            self.epoch += used
            self.steps_done = cfg.start_step + steps
            self._check_faults()
            return RunResult(self.steps_done, self.steps_done, False, float("nan"), ev0.elapsed_time(ev1) / 1e3,
                             self.kernel_launches - l0, 0, cfg.nx, cfg.ny, steps)
        if len(self.engines) != 1 or self.world != 1 or cfg.backend not in ("fused", "single"):
            raise ValueError("solve_pipelined needs the fused backend on a single tile or a 1 x P grid")
        plan = self.engines[0].plan
        l0 = plan.launch_count
        ev0 = torch.cuda.Event(enable_timing=True)
        ev1 = torch.cuda.Event(enable_timing=True)
        with torch.cuda.device(self.device):
            ev0.record()
            self.parity = plan.solve_pipelined(host_in, host_out, steps, cfg.k, block_rows, skew_rows)
            ev1.record()
            ev1.synchronize()
        self.steps_done = cfg.start_step + steps
        return RunResult(self.steps_done, self.steps_done, False, float("nan"), ev0.elapsed_time(ev1) / 1e3,
                         plan.launch_count - l0, 0, cfg.nx, cfg.ny, steps)

    def _run_plain(self, steps: int) -> RunResult:
        """Plain one-step kernel + standalone norm kernel, host-side decision (the
        structure of the reference loop; baseline for the fused path)."""
        cfg = self.cfg
        f = self.fields[0]
        ev0 = torch.cuda.Event(enable_timing=True)
        ev1 = torch.cuda.Event(enable_timing=True)
        converged, last_norm, checks, launches, iterations = False, float("nan"), 0, 0, None
        done0 = self.steps_done
        with torch.cuda.device(self.device):
            ev0.record()
            for s in range(self.steps_done, self.steps_done + steps):
                K.step_plain(f.buf[self.parity], f.buf[1 - self.parity], f.tile, self.phys)
                launches += 1
                if cfg.conv and s % cfg.interval == 0:
                    v = K.norm_diff(f.buf[1 - self.parity], f.buf[self.parity], f.tile, cfg.norm)
                    last_norm = float(v.item())
                    launches += 1
                    checks += 1
                    if last_norm < cfg.eps:
                        converged, iterations = True, s
                        break
                self.parity ^= 1
                self.steps_done += 1
            ev1.record()
            ev1.synchronize()
        if iterations is None:
            iterations = self.steps_done
        return RunResult(self.steps_done, iterations, converged, last_norm, ev0.elapsed_time(ev1) / 1e3,
                         launches, checks, cfg.nx, cfg.ny, self.steps_done - done0)

    def _run_resident(self, steps: int) -> RunResult:
        cfg = self.cfg
        if cfg.conv:
            raise ValueError("the resident backend has no convergence test; use backend='fused'")
        f = self.fields[0]
        ev0 = torch.cuda.Event(enable_timing=True)
        ev1 = torch.cuda.Event(enable_timing=True)
        with torch.cuda.device(self.device):
            ev0.record()
            K.resident_run(f.owned(self.parity), f.owned(1 - self.parity), steps, self.phys)
            ev1.record()
            ev1.synchronize()
        self.parity ^= 1
        self.steps_done += steps
        return RunResult(self.steps_done, self.steps_done, False, float("nan"),
                         ev0.elapsed_time(ev1) / 1e3, 1, 0, cfg.nx, cfg.ny, steps)

    def _check_faults(self):
        for e in self.engines:
            code = e.fault()
            if code:
                raise RuntimeError(
                    f"rank {e.rank}: flag wait timed out (code {code}: " + (
                        "task dependency inside the dataflow launch" if code == 50 else
                        "direction " + N.DIRS[code - 1] if code < 50 else
                        "norm slot of rank " + str(code - 100)) + ")")

    # ----------------------------------------------------------------- output
    def gather(self, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Assemble the current field as one nx x ny host tensor (pinned if given)."""
        cfg = self.cfg
        if out is None:
            out = torch.empty((cfg.nx, cfg.ny), dtype=torch.float32)
        for f in self.fields:
            t = f.tile
            out[t.x0:t.x0 + t.rows, t.y0:t.y0 + t.cols].copy_(f.owned(self.parity), non_blocking=True)
        torch.cuda.synchronize(self.device)
        return out

    def local_tiles(self):
        """[(x0, y0, host ndarray)] of the tiles this process owns (the per-rank dump)."""
        out = []
        for f in self.fields:
            out.append((f.tile.x0, f.tile.y0, f.owned(self.parity).cpu().numpy()))
        return out

    def owned_views(self) -> List[torch.Tensor]:
        return [f.owned(self.parity) for f in self.fields]

    def halo_wait_stats(self):
        """Sum over this process's tiles of (us blocked on neighbour flags, blocking waits)."""
        tot = [e.halo_wait_stats() for e in self.engines]
        return sum(t[0] for t in tot), sum(t[1] for t in tot)

    @property
    def kernel_launches(self) -> int:
        return sum(e.plan.launch_count for e in self.engines)

    def close(self):
        for e in self.engines:
            if e.plan is not None:
                e.plan.close()
        self.engines = []
