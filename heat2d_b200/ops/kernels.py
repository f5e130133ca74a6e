"""Thin Python wrappers over the C ABI (tensors in, raw pointers out).

Every CUDA op launches one of *our* kernels on the current torch stream; there
is no PyTorch fallback — a missing ``libheat2d_cuda.so`` raises.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np

from .. import _native as N
from .._lazy import LazyModule

torch = LazyModule("torch")   # only the CUDA wrappers need it; the CPU oracle does not

# ------------------------------------------------------------------- helpers


def _stream_ptr(device=None) -> int:
    return int(torch.cuda.current_stream(device).cuda_stream)


def _chk_dev(t: torch.Tensor, name: str):
    if not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor")
    if t.dtype != torch.float32:
        raise ValueError(f"{name} must be float32")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous (padded array)")


def _padded_ok(t: torch.Tensor, tile: N.Tile, name: str):
    need = (tile.rows + 2 * tile.halo_r, tile.pitch)
    if tuple(t.shape) != need:
        raise ValueError(f"{name}: expected padded shape {need}, got {tuple(t.shape)}")


# ---------------------------------------------------------------------- CPU


def cpu_init(base: np.ndarray, tile: N.Tile, mode: str = "exact", with_margins: bool = True):
    """Fill a padded host array with the initial condition (inidat, mpi_heat2Dn.c:242-248)."""
    assert base.dtype == np.float32 and base.flags.c_contiguous
    lib = N.cpu()
    N.check(lib, lib.h2d_cpu_init(base.ctypes.data, C.byref(tile), N.INIT[mode], int(with_margins)))
    return base


def cpu_initial_grid(nx: int, ny: int, mode: str = "exact") -> np.ndarray:
    u = np.zeros((nx, ny), dtype=np.float32)
    cpu_init(u, N.make_tile(nx, ny), mode, False)
    return u


def cpu_run(u: np.ndarray, steps: int, phys: N.Phys, norm: str = "none", interval: int = 20,
            eps: float = 0.1, threads: int = 1):
    """Oracle: advance the dense grid `u` in place.  Returns (steps_done, last_norm)."""
    assert u.dtype == np.float32 and u.flags.c_contiguous and u.ndim == 2
    lib = N.cpu()
    last = C.c_double(float("nan"))
    done = lib.h2d_cpu_run(u.ctypes.data, u.shape[0], u.shape[1], int(steps), C.byref(phys),
                           N.NORM[norm], int(interval), float(eps), int(threads), C.byref(last))
    N.check(lib, done)
    return int(done), float(last.value)


def cpu_step_tile(src: np.ndarray, dst: np.ndarray, tile: N.Tile, phys: N.Phys):
    lib = N.cpu()
    N.check(lib, lib.h2d_cpu_step_tile(src.ctypes.data, dst.ctypes.data, C.byref(tile), C.byref(phys)))


# --------------------------------------------------------------------- CUDA


def device_info(device: int = 0) -> str:
    lib = N.cuda()
    buf = C.create_string_buffer(1024)
    N.check(lib, lib.h2d_cuda_device_info(device, buf, 1024))
    return buf.value.decode()


def init_field(base: torch.Tensor, tile: N.Tile, mode: str = "exact", with_margins: bool = True):
    """K1: initial condition on the device, margins included (so no start-up
    halo exchange is needed: every rank can evaluate the formula in its ghosts)."""
    _chk_dev(base, "base")
    _padded_ok(base, tile, "base")
    lib = N.cuda()
    with torch.cuda.device(base.device):
        N.check(lib, lib.h2d_cuda_init(base.data_ptr(), C.byref(tile), N.INIT[mode], int(with_margins),
                                       _stream_ptr(base.device)))
    return base


def step_plain(src: torch.Tensor, dst: torch.Tensor, tile: N.Tile, phys: N.Phys):
    """K2/K3: one time step with the plain (k = 1, no TMA) kernel."""
    _chk_dev(src, "src"); _chk_dev(dst, "dst")
    _padded_ok(src, tile, "src"); _padded_ok(dst, tile, "dst")
    lib = N.cuda()
    with torch.cuda.device(src.device):
        N.check(lib, lib.h2d_cuda_step_k1(src.data_ptr(), dst.data_ptr(), C.byref(tile), C.byref(phys),
                                          _stream_ptr(src.device)))


def norm_diff(a: torch.Tensor, b: torch.Tensor, tile: N.Tile, norm: str,
              scratch: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None):
    """norm of (a - b) over the owned cells -> 1-element device tensor."""
    _chk_dev(a, "a"); _chk_dev(b, "b")
    lib = N.cuda()
    need = lib.h2d_norm_scratch_bytes(C.byref(tile))
    if scratch is None:
        scratch = torch.zeros((need + 3) // 4, dtype=torch.int32, device=a.device)
    if out is None:
        out = torch.zeros(1, dtype=torch.float32, device=a.device)
    with torch.cuda.device(a.device):
        N.check(lib, lib.h2d_cuda_norm_diff(a.data_ptr(), b.data_ptr(), C.byref(tile), N.NORM[norm],
                                            scratch.data_ptr(), out.data_ptr(), _stream_ptr(a.device)))
    return out


def resident_fits(nx: int, ny: int) -> int:
    return int(N.cuda().h2d_resident_fits(nx, ny))


def resident_run(src: torch.Tensor, dst: torch.Tensor, steps: int, phys: N.Phys):
    """Whole run in one cluster launch (dense nx x ny tensors, any row pitch)."""
    assert src.is_cuda and dst.is_cuda and src.dtype == torch.float32 and dst.dtype == torch.float32
    assert src.dim() == 2 and src.shape == dst.shape and src.stride(1) == 1 and dst.stride(1) == 1
    assert src.stride(0) == dst.stride(0)
    lib = N.cuda()
    with torch.cuda.device(src.device):
        N.check(lib, lib.h2d_resident_run(src.data_ptr(), dst.data_ptr(), src.shape[0], src.shape[1],
                                          src.stride(0), int(steps), C.byref(phys),
                                          _stream_ptr(src.device)))


def pack_edge(base: torch.Tensor, tile: N.Tile, direction: str, w: int, out: torch.Tensor):
    lib = N.cuda()
    with torch.cuda.device(base.device):
        N.check(lib, lib.h2d_cuda_pack_edge(base.data_ptr(), C.byref(tile), N.DIRS.index(direction), w,
                                            out.data_ptr(), _stream_ptr(base.device)))
    return out


def unpack_ghost(base: torch.Tensor, tile: N.Tile, direction: str, w: int, inp: torch.Tensor):
    lib = N.cuda()
    with torch.cuda.device(base.device):
        N.check(lib, lib.h2d_cuda_unpack_ghost(base.data_ptr(), C.byref(tile), N.DIRS.index(direction), w,
                                               inp.data_ptr(), _stream_ptr(base.device)))


class PeerRef:
    """What a plan needs to know about one neighbour (all raw device addresses)."""

    def __init__(self, buf0: int, buf1: int, flag: int, x0: int, y0: int, halo_r: int, halo_c: int,
                 pitch: int, colbuf0: int = 0, colbuf1: int = 0):
        self.buf0, self.buf1, self.flag = int(buf0), int(buf1), int(flag)
        self.x0, self.y0, self.halo_r, self.halo_c, self.pitch = x0, y0, halo_r, halo_c, int(pitch)
        # W / E neighbours: their packed column-ghost buffers (parity 0 / 1) facing this tile
        self.colbuf0, self.colbuf1 = int(colbuf0), int(colbuf1)


class Plan:
    """The temporal-blocked, TMA-fed, halo-fusing stencil for one tile (K4-K7)."""

    def __init__(self, tile: N.Tile, phys: N.Phys, buf0_ptr: int, buf1_ptr: int, kmax: int,
                 chunk_rows: int = 0, peers: Optional[dict] = None, flags_in_ptr: int = 0,
                 scratch_ptr: int = 0, spin_timeout_s: float = 10.0, device=None, layout: int = 0,
                 colbuf_in=None, team: int = 0):
        self._lib = N.cuda()
        self.tile, self.phys, self.kmax = tile, phys, kmax
        self.device = device
        d = N.PlanDesc()
        d.tile = tile
        d.phys = phys
        d.buf[0] = buf0_ptr
        d.buf[1] = buf1_ptr
        d.kmax = kmax
        d.chunk_rows = chunk_rows
        if peers:
            for name, p in peers.items():
                if p is None:
                    continue
                e = d.peer[N.DIRS.index(name)]
                e.buf[0] = p.buf0
                e.buf[1] = p.buf1
                e.flag = p.flag
                e.x0, e.y0, e.halo_r, e.halo_c, e.pitch = p.x0, p.y0, p.halo_r, p.halo_c, p.pitch
                e.colbuf[0] = p.colbuf0 or None
                e.colbuf[1] = p.colbuf1 or None
        d.flags_in = flags_in_ptr or None
        d.scratch = scratch_ptr or None
        d.spin_timeout_ns = int(spin_timeout_s * 1e9)
        d.layout = int(layout)
        d.team = int(team)     # eligible sweeps as cluster launches (h2d_plan_desc_t::team)
        if colbuf_in is not None:      # [[west parity 0, west parity 1], [east parity 0, east parity 1]]
            for side in (0, 1):
                for par in (0, 1):
                    d.colbuf_in[side][par] = int(colbuf_in[side][par]) or None
        self._desc = d
        with torch.cuda.device(device):
            h = self._lib.h2d_plan_create(C.byref(d))
        if not h:
            raise N.NativeError(-1, N.last_error(self._lib))
        self._h = h

    @staticmethod
    def scratch_bytes(tile: N.Tile, kmax: int, chunk_rows: int = 0) -> int:
        return int(N.cuda().h2d_plan_scratch_bytes(C.byref(tile), kmax, chunk_rows))

    def launch(self, src_parity: int, k: int, epoch: int, norm: str = "none", ctl_ptr: int = 0,
               run_if: int = 0):
        """One k-step launch.  With `ctl_ptr` the launch is a device-side no-op
        unless the word at ctl_ptr equals `run_if` (see h2d_plan_launch_ex)."""
        with torch.cuda.device(self.device):
            N.check(self._lib, self._lib.h2d_plan_launch_ex(self._h, src_parity, k, epoch & 0xFFFFFFFF,
                                                            N.NORM[norm], ctl_ptr or None, run_if & 0xFFFFFFFF,
                                                            _stream_ptr(self.device)))

    @staticmethod
    def run_many(plans, src_parity: int, k: int, n_launches: int, epoch0: int, ctl_ptrs=None):
        """`n_launches` back-to-back k-step launches of every plan in `plans` (one
        native call: the per-launch cost is the bare cudaLaunchKernel).  `ctl_ptrs`
        (one device word per plan) puts them under device-side run control."""
        lib = plans[0]._lib
        arr = (C.c_void_p * len(plans))(*[p._h for p in plans])
        ctl = (C.c_void_p * len(plans))(*ctl_ptrs) if ctl_ptrs else None
        with torch.cuda.device(plans[0].device):
            N.check(lib, lib.h2d_plan_run_many(arr, len(plans), src_parity, k, n_launches,
                                               epoch0 & 0xFFFFFFFF, ctl, _stream_ptr(plans[0].device)))

    def run_dataflow(self, src_parity: int, k: int, n_epochs: int) -> int:
        """`n_epochs` k-step sweeps in one dataflow launch (h2d_plan_run_dataflow);
        returns the parity of the buffer that holds the result."""
        with torch.cuda.device(self.device):
            N.check(self._lib, self._lib.h2d_plan_run_dataflow(self._h, src_parity, k, n_epochs,
                                                                 _stream_ptr(self.device)))
        return src_parity ^ (n_epochs & 1)

    def launch_team(self, src_parity: int, k: int, cluster: int = 2, chunk_rows: int = 0) -> int:
        """One k-step sweep (5 <= k <= 8) as a cluster launch: teams of 4 * cluster adjacent strips
        that share their edge columns through (distributed) shared memory (h2d_plan_launch_team).
        Returns the parity of the buffer that holds the result."""
        with torch.cuda.device(self.device):
            N.check(self._lib, self._lib.h2d_plan_launch_team(self._h, src_parity, k, cluster, chunk_rows,
                                                                _stream_ptr(self.device)))
        return src_parity ^ 1

    @property
    def team_fallback_ptr(self) -> int:
        """Device address of the counter of team strips whose data left the fast strict evaluation."""
        return int(self._lib.h2d_plan_team_fallback_ptr(self._h))

    def solve_pipelined(self, host_in: torch.Tensor, host_out: torch.Tensor, steps: int, k: int,
                        block_rows: int = 0, skew_rows: int = 0) -> int:
        """Upload `host_in`, advance `steps` steps, download into `host_out`, overlapped
        (h2d_plan_solve_pipelined).  Both tensors: pinned, float32, (rows, cols), unit
        column stride.  Returns the parity of the buffer that holds the result."""
        for t_, nm in ((host_in, "host_in"), (host_out, "host_out")):
            if t_.is_cuda or t_.dtype != torch.float32 or t_.dim() != 2 or t_.stride(1) != 1:
                raise ValueError(f"{nm} must be a 2-D float32 host tensor with unit column stride")
            if not t_.is_pinned():
                raise ValueError(f"{nm} must be pinned (torch.Tensor.pin_memory) for the copies to overlap")
            if tuple(t_.shape) != (self.tile.rows, self.tile.cols):
                raise ValueError(f"{nm}: expected shape {(self.tile.rows, self.tile.cols)}, got {tuple(t_.shape)}")
        par = C.c_int(0)
        with torch.cuda.device(self.device):
            N.check(self._lib, self._lib.h2d_plan_solve_pipelined(
                self._h, host_in.data_ptr(), host_in.stride(0), host_out.data_ptr(), host_out.stride(0),
                int(steps), int(k), int(block_rows), int(skew_rows), _stream_ptr(self.device), C.byref(par)))
        return par.value

    @staticmethod
    def solve_pipelined_many(plans, host_in: torch.Tensor, host_out: torch.Tensor, col0s, steps: int, k: int,
                             epoch0: int, block_rows: int = 0, skew_rows: int = 0):
        """The pipelined solve for the tiles of a 1 x P grid emulated in one process: `host_in` /
        `host_out` are the WHOLE pinned nx x ny fields, tile i owns columns col0s[i] .. +cols (its
        ghost columns are simply the neighbouring columns of host_in).  Returns (parity, epochs)."""
        for t_, nm in ((host_in, "host_in"), (host_out, "host_out")):
            if t_.is_cuda or t_.dtype != torch.float32 or t_.dim() != 2 or t_.stride(1) != 1 or not t_.is_pinned():
                raise ValueError(f"{nm} must be a pinned 2-D float32 host tensor with unit column stride")
        n = len(plans)
        lib = plans[0]._lib
        arr = (C.c_void_p * n)(*[p._h for p in plans])
        pin = (C.c_void_p * n)(*[host_in.data_ptr() + 4 * int(c) for c in col0s])
        pout = (C.c_void_p * n)(*[host_out.data_ptr() + 4 * int(c) for c in col0s])
        ipitch = (C.c_int64 * n)(*([host_in.stride(0)] * n))
        opitch = (C.c_int64 * n)(*([host_out.stride(0)] * n))
        par, used = C.c_int(0), C.c_uint32(0)
        with torch.cuda.device(plans[0].device):
            N.check(lib, lib.h2d_plan_solve_pipelined_many(arr, n, pin, ipitch, pout, opitch, int(steps), int(k),
                                                           int(block_rows), int(skew_rows), epoch0 & 0xFFFFFFFF,
                                                           _stream_ptr(plans[0].device), C.byref(par),
                                                           C.byref(used)))
        return par.value, int(used.value)

    @property
    def wave_rows(self) -> int:
        """Rows one wave of window tasks covers (block granularity of the pipelined solve)."""
        return int(self._lib.h2d_plan_wave_rows(self._h))

    def solve_pipelined_strip(self, host_in: torch.Tensor, col0: int, host_out: torch.Tensor, steps: int, k: int,
                              epoch0: int, wave_rows: int, block_rows: int = 0, skew_rows: int = 0):
        """Pipelined upload / solve / download for one rank of a 1 x P column-strip run
        (h2d_plan_solve_pipelined_ex).  `host_in`: pinned (rows, >= col0 + cols + margin) tensor
        whose column `col0` is owned column 0 and which holds 32 valid ghost columns on each side
        that has a neighbour; `host_out`: pinned (rows, cols).  Returns (parity, epochs used)."""
        for t_, nm in ((host_in, "host_in"), (host_out, "host_out")):
            if t_.is_cuda or t_.dtype != torch.float32 or t_.dim() != 2 or t_.stride(1) != 1:
                raise ValueError(f"{nm} must be a 2-D float32 host tensor with unit column stride")
            if not t_.is_pinned():
                raise ValueError(f"{nm} must be pinned (torch.Tensor.pin_memory) for the copies to overlap")
        if host_in.shape[0] != self.tile.rows or tuple(host_out.shape) != (self.tile.rows, self.tile.cols):
            raise ValueError("host_in / host_out do not match the tile")
        if col0 < 0 or col0 + self.tile.cols > host_in.shape[1]:
            raise ValueError("col0 + cols exceeds host_in")
        par, used = C.c_int(0), C.c_uint32(0)
        with torch.cuda.device(self.device):
            N.check(self._lib, self._lib.h2d_plan_solve_pipelined_ex(
                self._h, host_in.data_ptr() + 4 * col0, host_in.stride(0), host_out.data_ptr(), host_out.stride(0),
                int(steps), int(k), int(block_rows), int(skew_rows), int(wave_rows), epoch0 & 0xFFFFFFFF,
                _stream_ptr(self.device), C.byref(par), C.byref(used)))
        return par.value, int(used.value)

    def seed_colbuf(self, parity: int):
        """After the padded array of `parity` was filled by anything but the neighbours' launches
        (initial condition, load, upload): copy its ghost columns into the packed buffers."""
        with torch.cuda.device(self.device):
            N.check(self._lib, self._lib.h2d_plan_seed_colbuf(self._h, int(parity), _stream_ptr(self.device)))

    @property
    def launch_count(self) -> int:
        return int(self._lib.h2d_plan_launch_count(self._h))

    @property
    def layout(self) -> str:
        """Lane layout the plan's kernels use: 'v4', 'ab2', 'ab4' or 'v8'."""
        return {1: "v4", 2: "ab2", 3: "ab4", 4: "v8"}.get(int(self._lib.h2d_plan_layout(self._h)), "?")

    @property
    def norm_ptr(self) -> int:
        return int(self._lib.h2d_plan_norm_ptr(self._h))

    @property
    def fault_ptr(self) -> int:
        return int(self._lib.h2d_plan_fault_ptr(self._h))

    @property
    def stats_ptr(self) -> int:
        return int(self._lib.h2d_plan_stats_ptr(self._h))

    @property
    def strict_fallback_ptr(self) -> int:
        return int(self._lib.h2d_plan_strict_fallback_ptr(self._h))

    def close(self):
        if getattr(self, "_h", None):
            self._lib.h2d_plan_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
