"""One process per GPU: the fused multi-GPU solver and the NCCL-halo baseline.

``torch.distributed`` is used for exactly three things: bootstrap (rank / world
from ``torchrun``), shipping 64-byte CUDA-IPC handles and tile geometry between
ranks at start-up, and reductions *outside* the step loop (max of the device
timers).  Inside the step loop of :class:`DistributedHeat2DSolver` there is no
NCCL and no host synchronisation: halos travel as plain stores from the
stencil kernel into the neighbour's mapped ghost margin, ordered by per-launch
epoch flags (csrc/stencil_tk.cu), and the convergence value is finished by a
one-block kernel over peer memory (csrc/peer.cu).

What this replaces in the reference (``grad1612_mpi_heat.c``): ``MPI_Init`` /
``Cart_create`` / ``Cart_shift`` (:42-44,74-81), the derived row/column types
(:140-144), 16 persistent requests + ``Startall``/``Waitall`` per step
(:210-235,244,274), ``MPI_Allreduce`` (:268), ``MPI_Barrier`` + ``MPI_Reduce(MAX)``
of the elapsed time (:206,280) and the MPI-IO dump (:178-190).

:class:`HaloExchangeSolver` is the baseline the fused path is measured against
("NCCL-halo": pack kernel -> grouped send/recv -> unpack -> plain stencil), and
doubles as the CPU (gloo) multi-process test vehicle for the topology code.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional

import numpy as np
import torch
import torch.distributed as dist

from .. import _native as N
from ..config import HeatConfig
from ..models.heat import (COL_PACK, Heat2DSolver, RunResult, TileEngine, TileField, _layout_code, padded_bytes,
                           padded_geometry, tensor_from_ptr)
from ..ops import kernels as K
from .decomp import DIRS, Decomposition, OPPOSITE


def init_distributed(backend: Optional[str] = None):
    """Join the process group described by torchrun's environment.

    Returns (rank, world, local_rank, device).  Single-process runs (no RANK in
    the environment) get a world of 1 without creating a group.
    """
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    use_cuda = torch.cuda.is_available()
    device = torch.device(f"cuda:{local}") if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(device)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29533")
        backend = backend or ("nccl" if use_cuda else "gloo")
        if backend == "nccl":
            dist.init_process_group(backend, rank=rank, world_size=world, device_id=device)
        else:
            dist.init_process_group(backend, rank=rank, world_size=world)
    return rank, world, local, device


def _align(v: int, a: int = 512) -> int:
    return (v + a - 1) // a * a


class PeerArena:
    """A cudaMalloc'ed slab holding everything the neighbours must reach:
    both padded tile buffers, the 8 incoming epoch flags and the reduction
    tables.  Exported as a CUDA-IPC handle, imported by every other rank."""

    def __init__(self, tile: N.Tile, world: int):
        self.lib = N.cuda()
        pb = padded_bytes(tile)
        self.off_buf = [0, _align(pb)]
        self.off_flags = self.off_buf[1] + _align(pb)
        self.off_vals = self.off_flags + 512
        self.off_nflags = self.off_vals + _align(2 * world * 4)
        # packed column-ghost buffers [side][parity], rows x 8 floats each
        cb = _align(tile.rows * COL_PACK * 4)
        base = self.off_nflags + _align(2 * world * 4)
        self.off_colbuf = [[base + (2 * side + par) * cb for par in (0, 1)] for side in (0, 1)]
        self.bytes = base + 4 * cb
        self.handle = self.lib.h2d_arena_create(self.bytes)
        if not self.handle:
            raise N.NativeError(-5, N.last_error(self.lib))
        self.ptr = int(self.lib.h2d_arena_ptr(self.handle))
        self.peers: Dict[int, int] = {}

    def export(self) -> bytes:
        buf = C.create_string_buffer(64)
        N.check(self.lib, self.lib.h2d_arena_export(self.handle, buf))
        return buf.raw

    def import_peer(self, rank: int, handle: bytes) -> int:
        p = self.lib.h2d_arena_import(handle)
        if not p:
            raise N.NativeError(-2, N.last_error(self.lib))
        self.peers[rank] = int(p)
        return int(p)

    def meta(self) -> dict:
        return {"off_buf": self.off_buf, "off_flags": self.off_flags, "off_vals": self.off_vals,
                "off_nflags": self.off_nflags, "off_colbuf": self.off_colbuf}

    def close(self):
        for p in self.peers.values():
            self.lib.h2d_arena_unmap(p)
        self.peers = {}
        if self.handle:
            self.lib.h2d_arena_destroy(self.handle)
            self.handle = None


class DistributedHeat2DSolver(Heat2DSolver):
    """Fused stencil + halo exchange over NVLink peer memory, one rank per GPU."""

    def __init__(self, cfg: HeatConfig, device=None):
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        wsize = dist.get_world_size() if dist.is_initialized() else 1
        cfg.validate(world_size=wsize)
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device())
        self.arena: Optional[PeerArena] = None
        self._wsize = wsize
        super().__init__(cfg, device)

    # only this rank's tile lives here
    def _build(self):
        cfg, dev = self.cfg, self.device
        w, r = self._wsize, self.rank
        if cfg.backend not in ("fused",):
            raise ValueError("DistributedHeat2DSolver implements backend='fused'")
        spec = self.decomp.tile(r)
        tile = padded_geometry(spec, cfg.nx, cfg.ny, cfg.halo)
        with torch.cuda.device(dev):
            self.arena = PeerArena(tile, w)
        a = self.arena
        shape = (spec.rows + 2 * cfg.halo, tile.pitch)
        bufs = [tensor_from_ptr(a.ptr + a.off_buf[p], shape, torch.float32, dev, owner=a) for p in (0, 1)]
        colbuf = [[tensor_from_ptr(a.ptr + a.off_colbuf[side][p], (spec.rows, COL_PACK), torch.float32, dev, owner=a)
                   for p in (0, 1)] for side in (0, 1)]
        f = TileField(spec, cfg.nx, cfg.ny, cfg.halo, dev, storage=bufs, colbuf=colbuf)
        self.fields = [f]
        flags_in = tensor_from_ptr(a.ptr + a.off_flags, (64,), torch.int32, dev, owner=a)
        vals = tensor_from_ptr(a.ptr + a.off_vals, (2 * w,), torch.float32, dev, owner=a)
        nflags = tensor_from_ptr(a.ptr + a.off_nflags, (2 * w,), torch.int32, dev, owner=a)
        eng = TileEngine(f, self.phys, cfg.k, cfg.chunk_rows, w, r, flags_in, vals, nflags,
                         cfg.spin_timeout_s, _layout_code(cfg), int(cfg.team == "on"))
        self.engines = [eng]

        # exchange handles + geometry (start-up only)
        mine = {"handle": a.export(), "tile": [getattr(tile, n) for n, _ in N.Tile._fields_], **a.meta()}
        metas: List[dict] = [None] * w
        if w > 1:
            dist.all_gather_object(metas, mine)
        else:
            metas = [mine]
        base: Dict[int, int] = {r: a.ptr}
        for q in range(w):
            if q != r:
                with torch.cuda.device(dev):
                    base[q] = a.import_peer(q, metas[q]["handle"])
        peers: Dict[str, Optional[K.PeerRef]] = {}
        for d, nb in spec.neighbours.items():
            if nb is None:
                peers[d] = None
                continue
            m = metas[nb]
            nt = dict(zip([n for n, _ in N.Tile._fields_], m["tile"]))
            slot = DIRS.index(OPPOSITE[d])
            # my W band lands in the west neighbour's EAST packed buffer and vice versa
            side = {"W": 1, "E": 0}.get(d)
            cb = [base[nb] + m["off_colbuf"][side][p] for p in (0, 1)] if side is not None else [0, 0]
            peers[d] = K.PeerRef(base[nb] + m["off_buf"][0], base[nb] + m["off_buf"][1],
                                 base[nb] + m["off_flags"] + 4 * slot, nt["x0"], nt["y0"], nt["halo_r"],
                                 nt["halo_c"], nt["pitch"], cb[0], cb[1])
        all_vals = [base[q] + metas[q]["off_vals"] for q in range(w)]
        all_nfl = [base[q] + metas[q]["off_nflags"] for q in range(w)]
        eng.build(peers if w > 1 else {}, all_vals, all_nfl)
        self.barrier()

    def barrier(self):
        torch.cuda.synchronize(self.device)
        if self._wsize > 1:
            dist.barrier()

    def initialize(self, mode: Optional[str] = None):
        super().initialize(mode)
        # nobody may store into my ghost margins before my buffers are initialised
        self.barrier()
        return self

    def load(self, grid, parity: int = 0):
        super().load(grid, parity)
        self.barrier()
        return self

    def load_tile(self, padded_host: torch.Tensor, parity: int = 0):
        """H2D of this rank's padded tile (ghosts included) from (pinned) host memory."""
        f = self.fields[0]
        f.buf[parity].copy_(padded_host, non_blocking=True)
        self.parity = parity
        self._seed_colbufs()
        self.barrier()
        return self

    def device_barrier(self):
        """Align the ranks' streams on the device (see TileEngine.device_barrier)."""
        if self._wsize > 1:
            self.norm_seq += 1
            self.engines[0].device_barrier(self.norm_seq)

    def run(self, steps: Optional[int] = None) -> RunResult:
        self.barrier()
        # the host barrier leaves tens of microseconds of launch skew between the ranks; the
        # device-side one puts every rank's start event within the NVLink latency of the others
        self.device_barrier()
        res = super().run(steps)
        if self._wsize > 1:
            # elapsed = max over ranks (MPI_Reduce(MAX), grad1612_mpi_heat.c:280)
            t = torch.tensor([res.elapsed_s], dtype=torch.float64, device=self.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            res.elapsed_s = float(t.item())
        return res

    GHOST_COLS = 32   # ghost columns the pipelined solve uploads per side (TK_WE_COLS)

    def host_strip(self, pinned: bool = True):
        """(host_in, col0, host_out) for `solve_pipelined`: a (rows, cols + ghosts) input whose
        column `col0` is this rank's owned column 0, and a (rows, cols) output."""
        spec = self.decomp.tile(self.rank)
        gw = self.GHOST_COLS if spec.neighbours["W"] is not None else 0
        ge = self.GHOST_COLS if spec.neighbours["E"] is not None else 0
        hin = torch.zeros((spec.rows, gw + spec.cols + ge), dtype=torch.float32, pin_memory=pinned)
        hout = torch.empty((spec.rows, spec.cols), dtype=torch.float32, pin_memory=pinned)
        return hin, gw, hout

    def solve_pipelined(self, host_in: torch.Tensor, col0: int, host_out: torch.Tensor,
                        steps: Optional[int] = None) -> RunResult:
        """End to end on a 1 x P process grid: every rank uploads its column strip (plus the
        ghost columns of the initial field), advances `steps` steps and downloads its result, the
        three phases overlapped block by block and the ranks in lockstep (see
        h2d_plan_solve_pipelined_ex).  `elapsed_s` = device time of the slowest rank."""
        cfg = self.cfg
        steps = cfg.steps if steps is None else steps
        if cfg.px != 1:
            raise ValueError("solve_pipelined needs column strips (px = 1): row strips serialise the ranks")
        if cfg.conv:
            raise ValueError("solve_pipelined has no convergence test; use run()")
        plan = self.engines[0].plan
        wave = torch.tensor([plan.wave_rows], dtype=torch.int64, device=self.device)
        if self._wsize > 1:
            dist.all_reduce(wave, op=dist.ReduceOp.MIN)
        l0 = plan.launch_count
        self.barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.device(self.device):
            ev0.record()
            self.parity, used = plan.solve_pipelined_strip(host_in, col0, host_out, steps, cfg.k, self.epoch + 1,
                                                           int(wave.item()))
            ev1.record()
            ev1.synchronize()
        self.epoch += used
        self.steps_done = cfg.start_step + steps
        el = ev0.elapsed_time(ev1) / 1e3
        if self._wsize > 1:
            t = torch.tensor([el], dtype=torch.float64, device=self.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            el = float(t.item())
        self._check_faults()
        return RunResult(self.steps_done, self.steps_done, False, float("nan"), el, plan.launch_count - l0, 0,
                         cfg.nx, cfg.ny, steps)

    def gather(self, out: Optional[torch.Tensor] = None, dst: int = 0) -> Optional[torch.Tensor]:
        """Assemble the global field on rank `dst` (device copies over NCCL)."""
        cfg = self.cfg
        f = self.fields[0]
        mine = f.owned(self.parity).contiguous()
        if self._wsize == 1:
            return mine.cpu() if out is None else out.copy_(mine)
        if self.rank == dst:
            if out is None:
                out = torch.empty((cfg.nx, cfg.ny), dtype=torch.float32)
            for q, spec in enumerate(self.decomp.tiles):
                if q == dst:
                    part = mine
                else:
                    part = torch.empty((spec.rows, spec.cols), dtype=torch.float32, device=self.device)
                    dist.recv(part, src=q)
                out[spec.x0:spec.x0 + spec.rows, spec.y0:spec.y0 + spec.cols].copy_(part)
            torch.cuda.synchronize(self.device)
            return out
        dist.send(mine, dst=dst)
        return None

    def close(self):
        super().close()
        if self._wsize > 1:
            self.barrier()
        if self.arena is not None:
            self.arena.close()
            self.arena = None


class HaloExchangeSolver:
    """The "NCCL-halo" BASELINE (never the product path): every step packs the
    four owned edges, exchanges them with grouped ``isend``/``irecv``, unpacks
    into the ghost margins and runs the plain one-step kernel — the same
    algorithm as ``grad1612_mpi_heat.c:231-275`` without the overlap.

    Works on CUDA (nccl, our pack/unpack/step kernels) and on CPU (gloo, oracle
    tile step) so that the topology code is testable without GPUs.
    """

    def __init__(self, cfg: HeatConfig, device=None):
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.cfg = cfg.validate(world_size=self.world)
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() \
                else torch.device("cpu")
        self.device = torch.device(device)
        self.on_gpu = self.device.type == "cuda"
        self.decomp = Decomposition(cfg.nx, cfg.ny, cfg.px, cfg.py, cfg.uneven)
        self.spec = self.decomp.tile(self.rank)
        self.phys = N.make_phys(cfg.cx_eff, cfg.cy_eff, cfg.boundary, cfg.arith)
        self.field = TileField(self.spec, cfg.nx, cfg.ny, max(cfg.halo, cfg.k, 1), self.device)
        self.parity = 0
        self.steps_done = cfg.start_step
        t = self.field.tile
        self._send = {d: torch.empty(t.cols if d in "NS" else t.rows, dtype=torch.float32, device=self.device)
                      for d in "NSWE"}
        self._recv = {d: torch.empty_like(self._send[d]) for d in "NSWE"}
        self.kernel_launches = 0
        # k > 1 (GPU only): the SAME temporal-blocked kernel as the fused product, without peers,
        # fed by one grouped NCCL exchange of k-wide bands per launch — the comparison that
        # isolates the halo mechanism (peer stores + flags vs pack / send / recv / unpack)
        self.plan = None
        if cfg.k > 1:
            if not self.on_gpu:
                raise ValueError("the nccl-halo baseline runs k > 1 on GPUs only (CPU / gloo: k = 1)")
            if cfg.conv:
                raise ValueError("the nccl-halo baseline checks convergence with k = 1 only")
            f = self.field
            self.plan = K.Plan(f.tile, self.phys, f.buf[0].data_ptr(), f.buf[1].data_ptr(), cfg.k,
                               cfg.chunk_rows, device=self.device, layout=_layout_code(cfg))

    def initialize(self, mode: Optional[str] = None):
        mode = mode or self.cfg.init
        f = self.field
        f.buf[1].zero_()
        if self.on_gpu:
            K.init_field(f.buf[0], f.tile, mode, True)
        else:
            K.cpu_init(f.buf[0].numpy(), f.tile, mode, True)
        self.parity = 0
        return self

    def load(self, grid, parity: int = 0):
        if isinstance(grid, np.ndarray):
            grid = torch.from_numpy(grid)
        self.field.buf[1 - parity].zero_()
        self.field.load_from_global(grid, parity)
        self.parity = parity
        return self

    def _exchange(self):
        f, t = self.field, self.field.tile
        src = f.buf[self.parity]
        ops = []
        nbs = self.spec.neighbours
        for d in "NSWE":
            if nbs[d] is None:
                continue
            if self.on_gpu:
                K.pack_edge(src, t, d, 1, self._send[d])
                self.kernel_launches += 1
            else:
                o = f.owned(self.parity)
                edge = {"N": o[0, :], "S": o[-1, :], "W": o[:, 0], "E": o[:, -1]}[d]
                self._send[d].copy_(edge)
            ops.append(dist.P2POp(dist.isend, self._send[d], nbs[d]))
            ops.append(dist.P2POp(dist.irecv, self._recv[d], nbs[d]))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        for d in "NSWE":
            if nbs[d] is None:
                continue
            if self.on_gpu:
                K.unpack_ghost(src, t, d, 1, self._recv[d])
                self.kernel_launches += 1
            else:
                hr, hc = t.halo_r, t.halo_c
                if d == "N":
                    src[hr - 1, hc:hc + t.cols] = self._recv[d]
                elif d == "S":
                    src[hr + t.rows, hc:hc + t.cols] = self._recv[d]
                elif d == "W":
                    src[hr:hr + t.rows, hc - 1] = self._recv[d]
                else:
                    src[hr:hr + t.rows, hc + t.cols] = self._recv[d]

    # ---- CPU path with the reference's overlap (grad1612_mpi_heat.c:233-259): start the halo
    # transfers, sweep the cells that need no ghost, wait, sweep the edge ring.
    def _sub_step(self, i0: int, i1: int, j0: int, j1: int):
        """One step on owned cells [i0, i1) x [j0, j1) of the tile (CPU arrays)."""
        if i1 <= i0 or j1 <= j0:
            return
        f, t = self.field, self.field.tile
        sub = N.Tile(t.nx, t.ny, t.x0 + i0, t.y0 + j0, i1 - i0, j1 - j0, t.halo_r, t.halo_c, t.pitch)
        off = i0 * t.pitch + j0
        src = f.buf[self.parity].numpy().reshape(-1)[off:]
        dst = f.buf[1 - self.parity].numpy().reshape(-1)[off:]
        K.cpu_step_tile(src, dst, sub, self.phys)

    def _exchange_and_step_overlapped(self):
        f, t = self.field, self.field.tile
        nbs = self.spec.neighbours
        o = f.owned(self.parity)
        ops = []
        for d in "NSWE":
            if nbs[d] is None:
                continue
            edge = {"N": o[0, :], "S": o[-1, :], "W": o[:, 0], "E": o[:, -1]}[d]
            self._send[d].copy_(edge)
            ops.append(dist.P2POp(dist.isend, self._send[d], nbs[d]))
            ops.append(dist.P2POp(dist.irecv, self._recv[d], nbs[d]))
        works = dist.batch_isend_irecv(ops) if ops else []
        self._sub_step(1, t.rows - 1, 1, t.cols - 1)            # interior: no ghost needed
        for w in works:
            w.wait()
        src = f.buf[self.parity]
        hr, hc = t.halo_r, t.halo_c
        for d in "NSWE":
            if nbs[d] is None:
                continue
            if d == "N":
                src[hr - 1, hc:hc + t.cols] = self._recv[d]
            elif d == "S":
                src[hr + t.rows, hc:hc + t.cols] = self._recv[d]
            elif d == "W":
                src[hr:hr + t.rows, hc - 1] = self._recv[d]
            else:
                src[hr:hr + t.rows, hc + t.cols] = self._recv[d]
        self._sub_step(0, 1, 0, t.cols)                         # edge rows ...
        self._sub_step(t.rows - 1, t.rows, 0, t.cols)
        self._sub_step(1, t.rows - 1, 0, 1)                     # ... and edge columns
        self._sub_step(1, t.rows - 1, t.cols - 1, t.cols)

    def _exchange_bands(self, w: int):
        """k-wide halo exchange in two grouped rounds: W/E column bands first, then N/S row bands
        INCLUDING the ghost columns just received, which carries the corners along (temporal
        blocking needs them; the reference, one cell per step, never did: grad1612_mpi_heat.c:140-144)."""
        f, t = self.field, self.field.tile
        src = f.buf[self.parity]
        hr, hc, R, Cn = t.halo_r, t.halo_c, t.rows, t.cols
        nbs = self.spec.neighbours
        for phase in ("WE", "NS"):
            ops, pending = [], []
            for d in phase:
                if nbs[d] is None:
                    continue
                if d == "W":
                    out, dst = src[hr:hr + R, hc:hc + w], src[hr:hr + R, hc - w:hc]
                elif d == "E":
                    out, dst = src[hr:hr + R, hc + Cn - w:hc + Cn], src[hr:hr + R, hc + Cn:hc + Cn + w]
                elif d == "N":
                    out, dst = src[hr:hr + w, hc - w:hc + Cn + w], src[hr - w:hr, hc - w:hc + Cn + w]
                else:
                    out, dst = src[hr + R - w:hr + R, hc - w:hc + Cn + w], src[hr + R:hr + R + w, hc - w:hc + Cn + w]
                sbuf = out.contiguous()
                rbuf = torch.empty_like(sbuf)
                ops.append(dist.P2POp(dist.isend, sbuf, nbs[d]))
                ops.append(dist.P2POp(dist.irecv, rbuf, nbs[d]))
                pending.append((dst, rbuf))
            if ops:
                for wk in dist.batch_isend_irecv(ops):
                    wk.wait()
            for dst, rbuf in pending:
                dst.copy_(rbuf)

    def _run_blocked(self, steps: int):
        """k steps per exchange: grouped NCCL exchange of k-wide bands, then one launch of the
        peer-less temporal-blocked kernel."""
        from ..config import make_schedule
        for L in make_schedule(steps, self.cfg.k, False, 20, "exact", self.steps_done):
            if self.world > 1:
                self._exchange_bands(L.k)
            self.plan.launch(self.parity, L.k, 0, "none")
            self.kernel_launches += 1
            self.parity ^= 1
            self.steps_done += L.k

    def _step(self):
        f = self.field
        if self.on_gpu:
            K.step_plain(f.buf[self.parity], f.buf[1 - self.parity], f.tile, self.phys)
            self.kernel_launches += 1
        else:
            K.cpu_step_tile(f.buf[self.parity].numpy(), f.buf[1 - self.parity].numpy(), f.tile, self.phys)

    def _norm(self) -> float:
        cfg, f = self.cfg, self.field
        if self.on_gpu:
            v = K.norm_diff(f.buf[1 - self.parity], f.buf[self.parity], f.tile, cfg.norm)
            self.kernel_launches += 1
        else:
            d = (f.owned(1 - self.parity).double() - f.owned(self.parity).double())
            v = (d * d).sum().float().reshape(1) if cfg.norm == "l2sq" else d.abs().max().float().reshape(1)
        if self.world > 1:
            dist.all_reduce(v, op=dist.ReduceOp.SUM if cfg.norm == "l2sq" else dist.ReduceOp.MAX)
        return float(v.item())

    def run(self, steps: Optional[int] = None) -> RunResult:
        import time
        cfg = self.cfg
        steps = cfg.steps if steps is None else steps
        if self.world > 1:
            if self.on_gpu:
                torch.cuda.synchronize(self.device)
            dist.barrier()
        if self.on_gpu:
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
        t0 = time.perf_counter()
        converged, last, checks, iterations = False, float("nan"), 0, None
        done0 = self.steps_done
        if self.plan is not None:
            self._run_blocked(steps)
            steps = 0
        for s in range(self.steps_done, self.steps_done + steps):
            t_ = self.field.tile
            if self.world > 1 and not self.on_gpu and t_.rows >= 3 and t_.cols >= 3:
                self._exchange_and_step_overlapped()
            else:
                if self.world > 1:
                    self._exchange()
                self._step()
            if cfg.conv and s % cfg.interval == 0:
                last = self._norm()
                checks += 1
                if last < cfg.eps:
                    converged, iterations = True, s
                    break
            self.parity ^= 1
            self.steps_done += 1
        if self.on_gpu:
            ev1.record()
            ev1.synchronize()
            el = ev0.elapsed_time(ev1) / 1e3
        else:
            el = time.perf_counter() - t0
        if self.world > 1:
            t = torch.tensor([el], dtype=torch.float64, device=self.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            el = float(t.item())
        if iterations is None:
            iterations = self.steps_done
        return RunResult(self.steps_done, iterations, converged, last, el, self.kernel_launches, checks,
                         cfg.nx, cfg.ny, self.steps_done - done0)

    def gather(self, out: Optional[torch.Tensor] = None, dst: int = 0):
        cfg = self.cfg
        mine = self.field.owned(self.parity).contiguous()
        if self.world == 1:
            return mine.cpu()
        if self.rank == dst:
            if out is None:
                out = torch.empty((cfg.nx, cfg.ny), dtype=torch.float32)
            for q, spec in enumerate(self.decomp.tiles):
                if q == dst:
                    part = mine
                else:
                    part = torch.empty((spec.rows, spec.cols), dtype=torch.float32, device=self.device)
                    dist.recv(part, src=q)
                out[spec.x0:spec.x0 + spec.rows, spec.y0:spec.y0 + spec.cols].copy_(part)
            return out
        dist.send(mine, dst=dst)
        return None

    def local_tiles(self):
        """[(x0, y0, host ndarray)] of the tile this rank owns (the per-rank dump)."""
        t = self.field.tile
        return [(t.x0, t.y0, self.field.owned(self.parity).cpu().numpy())]

    def close(self):
        if self.plan is not None:
            self.plan.close()
            self.plan = None
