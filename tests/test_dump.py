"""Dump writers, byte-compared with a Python restatement of the reference's
fprintf loops (mpi_heat2Dn.c:253-268, grad1612_mpi_heat.c:195-200)."""
import numpy as np

from heat2d_b200.utils import dump as D
from refimpl import np_init


def py_prtdat(u):
    nx, ny = u.shape
    out = []
    for y in range(ny - 1, -1, -1):
        for x in range(nx):
            out.append("%6.1f" % float(u[x, y]))
            out.append(" " if x != nx - 1 else "\n")
    return "".join(out)


def py_rowmajor(u):
    nx, ny = u.shape
    return "".join("".join("%6.1f " % float(u[x, y]) for y in range(ny)) + "\n" for x in range(nx))


def test_prtdat_bytes(tmp_path):
    u = np_init(10, 12, "exact")
    u[3, 4] = -1234567.25
    u[5, 5] = 0.05
    p = tmp_path / "initial.dat"
    D.write_prtdat(str(p), u)
    assert p.read_text() == py_prtdat(u)


def test_rowmajor_bytes(tmp_path):
    u = np_init(9, 7, "exact") * np.float32(0.37)
    p = tmp_path / "final.dat"
    D.write_rowmajor(str(p), u)
    assert p.read_text() == py_rowmajor(u)


def test_large_values_and_negative_zero(tmp_path):
    u = np.array([[1.15e18, -0.0, 3.4e38], [1e-3, -0.04, 99999.95]], dtype=np.float32)
    p = tmp_path / "x.dat"
    D.write_rowmajor(str(p), u)
    assert p.read_text() == py_rowmajor(u)


def test_binary_tiles_assemble_and_convert(tmp_path):
    nx, ny = 12, 10
    u = np_init(nx, ny, "exact")
    p = tmp_path / "final_binary.dat"
    # four "ranks" write their sub-arrays, as the MPI-IO subarray view did
    for x0, rows in ((0, 7), (7, 5)):
        for y0, cols in ((0, 4), (4, 6)):
            D.write_binary_tile(str(p), u[x0:x0 + rows, y0:y0 + cols], nx, ny, x0, y0)
    assert p.stat().st_size == nx * ny * 4
    assert np.array_equal(D.read_binary(str(p), nx, ny), u)
    t = tmp_path / "final.dat"
    D.binary_to_text(str(p), str(t), nx, ny, "rowmajor")
    assert t.read_text() == py_rowmajor(u)
    D.binary_to_text(str(p), str(t), nx, ny, "prtdat")
    assert t.read_text() == py_prtdat(u)


def test_binary_keeps_stale_tail_unless_truncated(tmp_path):
    # grad1612_mpi_heat.c:178 opens WRONLY|CREATE without truncation
    p = tmp_path / "b.dat"
    big = np.ones((8, 8), dtype=np.float32)
    D.write_binary(str(p), big, truncate=True)
    small = np.zeros((4, 4), dtype=np.float32)
    D.write_binary_tile(str(p), small, 4, 4, 0, 0, truncate=False)
    assert p.stat().st_size == 8 * 8 * 4
    D.write_binary(str(p), small, truncate=True)
    assert p.stat().st_size == 4 * 4 * 4


def test_dump_dispatch(tmp_path):
    u = np_init(6, 5, "exact")
    assert D.dump("none", str(tmp_path), "initial", u) is None
    assert D.dump("prtdat", str(tmp_path), "initial", u).endswith("initial.dat")
    assert D.dump("binary", str(tmp_path), "final", u).endswith("final_binary.dat")


def test_own_6_1f_formatter_equals_printf():
    """The text writers format "%6.1f" themselves (csrc/dump.cpp::fmt61); every sampled float
    bit pattern — ties, negative zero, subnormals, huge values, inf, nan — must give printf's bytes.
    (All 2^32 patterns were compared once when the formatter was written: no mismatch.)"""
    import ctypes as C
    from heat2d_b200 import _native as N
    lib = N.cpu()
    lib.h2d_format_6_1f_check.argtypes = [C.c_uint32, C.c_uint64, C.c_uint32, C.POINTER(C.c_uint32)]
    lib.h2d_format_6_1f_check.restype = C.c_int64
    lib.h2d_format_6_1f.argtypes = [C.c_float, C.c_char_p]
    bad = C.c_uint32(0)
    # 4 Mi patterns spread over the whole 32-bit space, plus a dense run around 1.0 and around -0.05
    assert lib.h2d_format_6_1f_check(0, 1 << 22, 1021, C.byref(bad)) == 0, hex(bad.value)
    assert lib.h2d_format_6_1f_check(0x3F7FF000, 1 << 16, 1, C.byref(bad)) == 0, hex(bad.value)
    assert lib.h2d_format_6_1f_check(0xBD4CC000, 1 << 16, 1, C.byref(bad)) == 0, hex(bad.value)
    buf = C.create_string_buffer(64)
    for v in (0.0, -0.0, 0.05, 0.25, 0.75, -0.04, 9.95, 99999.95, 1e18, 3e38, float("inf"), float("nan")):
        n = lib.h2d_format_6_1f(v, buf)
        assert buf.raw[:n].decode() == "%6.1f" % C.c_float(v).value
