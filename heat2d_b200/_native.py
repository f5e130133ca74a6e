"""ctypes bindings of the two native libraries (see include/heat2d.h).

The libraries are plain C-ABI shared objects built in-tree by
:mod:`heat2d_b200.build`; nothing here depends on a torch C++ extension, so the
same ``.so`` files serve C callers.  On a box with a GPU the
CUDA library is mandatory: if it cannot be loaded the ops raise instead of
silently falling back to PyTorch.
"""
from __future__ import annotations

import ctypes as C
import threading
from pathlib import Path

from . import build as _build

# ---------------------------------------------------------------- constants
BOUNDARY = {"fixed": 0, "ghost-ring": 1}
ARITH = {"fp32": 0, "strict": 1, "strict-fma": 2}
INIT = {"ref-int32": 0, "ref-uint32": 1, "exact": 2, "zero": 3}
NORM = {"none": 0, "off": 0, "l2sq": 1, "linf": 2}
DIRS = ("N", "S", "W", "E", "NW", "NE", "SW", "SE")
NDIR = 8
KMAX = 8
NORM_STATIC_MAX_ROWS = 131072

H2D_OK, H2D_EINVAL, H2D_ECUDA, H2D_EIO, H2D_ETIMEOUT, H2D_ENOMEM = 0, -1, -2, -3, -4, -5


class Tile(C.Structure):
    _fields_ = [
        ("nx", C.c_int32), ("ny", C.c_int32),
        ("x0", C.c_int32), ("y0", C.c_int32),
        ("rows", C.c_int32), ("cols", C.c_int32),
        ("halo_r", C.c_int32), ("halo_c", C.c_int32),
        ("pitch", C.c_int64),
    ]


class Phys(C.Structure):
    _fields_ = [("cx", C.c_double), ("cy", C.c_double), ("boundary", C.c_int32), ("arith", C.c_int32)]


class Peer(C.Structure):
    _fields_ = [
        ("buf", C.c_void_p * 2),
        ("flag", C.c_void_p),
        ("x0", C.c_int32), ("y0", C.c_int32),
        ("halo_r", C.c_int32), ("halo_c", C.c_int32),
        ("pitch", C.c_int64),
        ("colbuf", C.c_void_p * 2),
    ]


class PlanDesc(C.Structure):
    _fields_ = [
        ("tile", Tile),
        ("phys", Phys),
        ("buf", C.c_void_p * 2),
        ("kmax", C.c_int32),
        ("chunk_rows", C.c_int32),
        ("peer", Peer * NDIR),
        ("flags_in", C.c_void_p),
        ("scratch", C.c_void_p),
        ("spin_timeout_ns", C.c_int64),
        ("layout", C.c_int32),
        ("team", C.c_int32),
        ("colbuf_in", (C.c_void_p * 2) * 2),
    ]


class NativeError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"[heat2d native error {code}] {msg}")
        self.code = code


_lock = threading.Lock()
_cpu = None
_cuda = None


def _declare_common(lib):
    lib.h2d_last_error.restype = C.c_char_p
    lib.h2d_version.restype = C.c_char_p


def _load(path: Path):
    return C.CDLL(str(path))


def cpu():
    """libheat2d_cpu.so (built on demand)."""
    global _cpu
    with _lock:
        if _cpu is None:
            path = _build.build_cpu()
            lib = _load(path)
            _declare_common(lib)
            P = C.POINTER
            lib.h2d_cpu_init.argtypes = [C.c_void_p, P(Tile), C.c_int, C.c_int]
            lib.h2d_cpu_run.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int64, P(Phys), C.c_int,
                                        C.c_int64, C.c_double, C.c_int, P(C.c_double)]
            lib.h2d_cpu_run.restype = C.c_int64
            lib.h2d_cpu_step_tile.argtypes = [C.c_void_p, C.c_void_p, P(Tile), P(Phys)]
            lib.h2d_dump_prtdat.argtypes = [C.c_char_p, C.c_void_p, C.c_int32, C.c_int32]
            lib.h2d_dump_rowmajor.argtypes = [C.c_char_p, C.c_void_p, C.c_int32, C.c_int32]
            lib.h2d_dump_binary_tile.argtypes = [C.c_char_p, C.c_void_p, C.c_int64, P(Tile), C.c_int]
            lib.h2d_binary_to_text.argtypes = [C.c_char_p, C.c_char_p, C.c_int32, C.c_int32, C.c_int]
            _cpu = lib
        return _cpu


def cuda():
    """libheat2d_cuda.so (built on demand).  Raises if it cannot be loaded."""
    global _cuda
    with _lock:
        if _cuda is None:
            path = _build.build_cuda()
            lib = _load(path)
            _declare_common(lib)
            P = C.POINTER
            vp = C.c_void_p
            lib.h2d_cuda_device_info.argtypes = [C.c_int, C.c_char_p, C.c_size_t]
            lib.h2d_cuda_init.argtypes = [vp, P(Tile), C.c_int, C.c_int, vp]
            lib.h2d_cuda_step_k1.argtypes = [vp, vp, P(Tile), P(Phys), vp]
            lib.h2d_plan_scratch_bytes.argtypes = [P(Tile), C.c_int32, C.c_int32]
            lib.h2d_plan_scratch_bytes.restype = C.c_size_t
            lib.h2d_plan_solve_pipelined_ex.argtypes = [vp, vp, C.c_int64, vp, C.c_int64, C.c_int64, C.c_int, C.c_int,
                                                        C.c_int, C.c_int, C.c_uint32, vp, P(C.c_int),
                                                        P(C.c_uint32)]
            lib.h2d_plan_solve_pipelined_ex.restype = C.c_int
            lib.h2d_plan_solve_pipelined_many.argtypes = [P(vp), C.c_int, P(vp), P(C.c_int64), P(vp), P(C.c_int64),
                                                          C.c_int64, C.c_int, C.c_int, C.c_int, C.c_uint32, vp,
                                                          P(C.c_int), P(C.c_uint32)]
            lib.h2d_plan_solve_pipelined_many.restype = C.c_int
            lib.h2d_plan_wave_rows.argtypes = [vp]
            lib.h2d_plan_wave_rows.restype = C.c_int32
            lib.h2d_plan_run_dataflow.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
            lib.h2d_plan_run_dataflow.restype = C.c_int
            lib.h2d_plan_launch_team.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
            lib.h2d_plan_launch_team.restype = C.c_int
            lib.h2d_plan_team_fallback_ptr.argtypes = [C.c_void_p]
            lib.h2d_plan_team_fallback_ptr.restype = C.c_void_p
            lib.h2d_plan_chunk_rows.argtypes = [P(Tile), C.c_int32, C.c_int32]
            lib.h2d_plan_chunk_rows.restype = C.c_int32
            lib.h2d_plan_chunk_rows_ex.argtypes = [P(Tile), C.c_int32, C.c_int32, C.c_int32, C.c_int32]
            lib.h2d_plan_chunk_rows_ex.restype = C.c_int32
            lib.h2d_plan_strip_width.argtypes = [C.c_int32, C.c_int32]
            lib.h2d_plan_strip_width.restype = C.c_int32
            lib.h2d_plan_layout.argtypes = [vp]
            lib.h2d_plan_layout.restype = C.c_int32
            lib.h2d_plan_seed_colbuf.argtypes = [vp, C.c_int, vp]
            lib.h2d_plan_create.argtypes = [P(PlanDesc)]
            lib.h2d_plan_create.restype = vp
            lib.h2d_plan_destroy.argtypes = [vp]
            lib.h2d_plan_destroy.restype = None
            lib.h2d_plan_launch.argtypes = [vp, C.c_int, C.c_int, C.c_uint32, C.c_int, vp]
            lib.h2d_plan_launch_ex.argtypes = [vp, C.c_int, C.c_int, C.c_uint32, C.c_int, vp, C.c_uint32, vp]
            lib.h2d_plan_run_many.argtypes = [C.POINTER(vp), C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint32,
                                              C.POINTER(vp), vp]
            lib.h2d_plan_solve_pipelined.argtypes = [vp, vp, C.c_int64, vp, C.c_int64, C.c_int64, C.c_int, C.c_int,
                                                     C.c_int, vp, C.POINTER(C.c_int)]
            lib.h2d_plan_launch_count.argtypes = [vp]
            lib.h2d_plan_launch_count.restype = C.c_int64
            lib.h2d_plan_norm_ptr.argtypes = [vp]
            lib.h2d_plan_norm_ptr.restype = vp
            lib.h2d_plan_fault_ptr.argtypes = [vp]
            lib.h2d_plan_fault_ptr.restype = vp
            lib.h2d_plan_stats_ptr.argtypes = [vp]
            lib.h2d_plan_stats_ptr.restype = vp
            lib.h2d_plan_strict_fallback_ptr.argtypes = [vp]
            lib.h2d_plan_strict_fallback_ptr.restype = vp
            lib.h2d_resident_fits.argtypes = [C.c_int32, C.c_int32]
            lib.h2d_resident_run.argtypes = [vp, vp, C.c_int32, C.c_int32, C.c_int64, C.c_int64, P(Phys), vp]
            lib.h2d_norm_scratch_bytes.argtypes = [P(Tile)]
            lib.h2d_norm_scratch_bytes.restype = C.c_size_t
            lib.h2d_cuda_norm_diff.argtypes = [vp, vp, P(Tile), C.c_int, vp, vp, vp]
            lib.h2d_cuda_norm_diff_static.argtypes = [vp, vp, P(Tile), C.c_int, vp, vp]
            lib.h2d_cuda_norm_finish.argtypes = [vp, vp, vp, C.c_int, C.c_int, C.c_uint32, C.c_int,
                                                 C.c_float, vp, vp, C.c_int, C.c_int64, vp, vp]
            lib.h2d_cuda_norm_finish_ctl.argtypes = [vp, vp, vp, C.c_int, C.c_int, C.c_uint32, C.c_int,
                                                     C.c_float, vp, vp, C.c_int, C.c_int64, vp, vp]
            lib.h2d_cuda_pack_edge.argtypes = [vp, P(Tile), C.c_int, C.c_int, vp, vp]
            lib.h2d_cuda_unpack_ghost.argtypes = [vp, P(Tile), C.c_int, C.c_int, vp, vp]
            lib.h2d_arena_create.argtypes = [C.c_size_t]
            lib.h2d_arena_create.restype = vp
            lib.h2d_arena_ptr.argtypes = [vp]
            lib.h2d_arena_ptr.restype = vp
            lib.h2d_arena_bytes.argtypes = [vp]
            lib.h2d_arena_bytes.restype = C.c_size_t
            lib.h2d_arena_export.argtypes = [vp, C.c_char_p]
            lib.h2d_arena_import.argtypes = [C.c_char_p]
            lib.h2d_arena_import.restype = vp
            lib.h2d_arena_unmap.argtypes = [vp]
            lib.h2d_arena_destroy.argtypes = [vp]
            lib.h2d_arena_destroy.restype = None
            _cuda = lib
        return _cuda


def check(lib, rc: int):
    """Raise :class:`NativeError` for a negative return code."""
    if rc is not None and rc < 0:
        raise NativeError(rc, lib.h2d_last_error().decode(errors="replace"))
    return rc


def last_error(lib) -> str:
    return lib.h2d_last_error().decode(errors="replace")


def make_tile(nx, ny, x0=0, y0=0, rows=None, cols=None, halo_r=0, halo_c=0, pitch=None) -> Tile:
    rows = nx if rows is None else rows
    cols = ny if cols is None else cols
    pitch = cols + 2 * halo_c if pitch is None else pitch
    return Tile(nx, ny, x0, y0, rows, cols, halo_r, halo_c, pitch)


def make_phys(cx=0.1, cy=0.1, boundary="fixed", arith="fp32") -> Phys:
    b = BOUNDARY[boundary] if isinstance(boundary, str) else int(boundary)
    a = ARITH[arith] if isinstance(arith, str) else int(arith)
    return Phys(float(cx), float(cy), b, a)
