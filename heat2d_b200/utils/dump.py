"""Grid dumps in the reference's three formats (native writers in csrc/dump.cpp).

* ``prtdat``    — ``mpi_heat2Dn.c:253-268`` (transposed, y descending)
* ``rowmajor``  — ``grad1612_mpi_heat.c:195-200`` (``"%6.1f "`` per value, newline per row)
* ``binary``    — raw fp32 row-major, the MPI-IO files ``initial_binary.dat`` /
  ``final_binary.dat`` (``grad1612_mpi_heat.c:178-190,283-285``); every rank
  ``pwrite``s its own rows into the shared file.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .. import _native as N


def _as_host(u) -> np.ndarray:
    # a torch tensor can only show up if torch is already imported: do not import it here
    # (the CPU backend runs without it)
    import sys
    torch = sys.modules.get("torch")
    if torch is not None and isinstance(u, torch.Tensor):
        u = u.detach().cpu().numpy()
    return np.ascontiguousarray(u, dtype=np.float32)


def write_prtdat(path: str, u):
    a = _as_host(u)
    lib = N.cpu()
    N.check(lib, lib.h2d_dump_prtdat(os.fsencode(path), a.ctypes.data, a.shape[0], a.shape[1]))


def write_rowmajor(path: str, u):
    a = _as_host(u)
    lib = N.cpu()
    N.check(lib, lib.h2d_dump_rowmajor(os.fsencode(path), a.ctypes.data, a.shape[0], a.shape[1]))


def write_binary_tile(path: str, owned, nx: int, ny: int, x0: int, y0: int, truncate: bool = False):
    """Write one tile's owned cells at their place in the shared nx x ny file."""
    a = _as_host(owned)
    t = N.make_tile(nx, ny, x0, y0, a.shape[0], a.shape[1])
    lib = N.cpu()
    N.check(lib, lib.h2d_dump_binary_tile(os.fsencode(path), a.ctypes.data, a.shape[1], C.byref(t),
                                          int(truncate)))


def write_binary(path: str, u, truncate: bool = True):
    a = _as_host(u)
    write_binary_tile(path, a, a.shape[0], a.shape[1], 0, 0, truncate)


def binary_to_text(bin_path: str, txt_path: str, nx: int, ny: int, fmt: str = "rowmajor"):
    lib = N.cpu()
    N.check(lib, lib.h2d_binary_to_text(os.fsencode(bin_path), os.fsencode(txt_path), nx, ny,
                                        1 if fmt == "prtdat" else 0))


def read_binary(path: str, nx: int, ny: int) -> np.ndarray:
    return np.fromfile(path, dtype=np.float32, count=nx * ny).reshape(nx, ny)


def dump(kind: str, out_dir: str, stem: str, u):
    """kind in {none, prtdat, rowmajor, binary}; stem in {initial, final}."""
    if kind == "none":
        return None
    os.makedirs(out_dir, exist_ok=True)
    if kind == "prtdat":
        p = os.path.join(out_dir, f"{stem}.dat")
        write_prtdat(p, u)
    elif kind == "rowmajor":
        p = os.path.join(out_dir, f"{stem}.dat")
        write_rowmajor(p, u)
    elif kind == "binary":
        p = os.path.join(out_dir, f"{stem}_binary.dat")
        write_binary(p, u)
    else:
        raise ValueError(kind)
    return p
