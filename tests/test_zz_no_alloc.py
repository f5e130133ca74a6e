"""The allocation-free norm entry point, called through the C ABI.

Some users of this library cannot allocate device memory at the point where
they need the convergence norm: the application's own pool has claimed the
whole GPU, or the call sits in a loop where cudaMalloc's device-wide
synchronisation is not acceptable.  For them h2d_cuda_norm_diff_static keeps
its scratch in a fixed array inside the library.  This test checks that it
gives the same answer as h2d_cuda_norm_diff (caller-provided scratch), twice in
a row — the second call reuses the array with the first call's contents still
in it — and that it refuses tiles beyond its documented limit.

Collected last on purpose (file name), so a run that stops at the first failure
has already exercised everything else.  It passes on its own in a fresh process
as well: it builds its inputs itself and relies on nothing the other tests did.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from heat2d_b200 import _native as N

pytestmark = pytest.mark.gpu


def _padded(nx, ny, hr, hc, seed, dev):
    pitch = (ny + 2 * hc + 31) // 32 * 32
    tile = N.make_tile(nx, ny, 0, 0, nx, ny, hr, hc, pitch)
    g = torch.Generator(device="cpu").manual_seed(seed)
    host = torch.zeros((nx + 2 * hr, pitch), dtype=torch.float32)
    host[hr:hr + nx, hc:hc + ny] = torch.rand((nx, ny), generator=g) * 1e3
    return tile, host.to(dev)


@pytest.mark.parametrize("norm", ["l2sq", "linf"])
def test_static_scratch_entry_point_matches_the_allocating_one(norm):
    lib = N.cuda()
    dev = torch.device("cuda:0")
    stream = int(torch.cuda.current_stream(dev).cuda_stream)
    code = N.NORM[norm]
    for call, (nx, ny) in enumerate([(1000, 777), (333, 2049)]):   # two calls in a row, different sizes
        tile, a = _padded(nx, ny, 8, 32, 10 + call, dev)
        _, b = _padded(nx, ny, 8, 32, 20 + call, dev)
        # existing entry point: scratch supplied by the caller
        need = lib.h2d_norm_scratch_bytes(C.byref(tile))
        scratch = torch.zeros((need + 3) // 4, dtype=torch.int32, device=dev)
        want = torch.zeros(1, dtype=torch.float32, device=dev)
        rc = lib.h2d_cuda_norm_diff(a.data_ptr(), b.data_ptr(), C.byref(tile), code, scratch.data_ptr(),
                                    want.data_ptr(), stream)
        assert rc == 0, N.last_error(lib)
        # new entry point: inputs and output only
        got = torch.full((1,), -1.0, dtype=torch.float32, device=dev)
        rc = lib.h2d_cuda_norm_diff_static(a.data_ptr(), b.data_ptr(), C.byref(tile), code, got.data_ptr(),
                                           stream)
        assert rc == 0, N.last_error(lib)
        torch.cuda.synchronize(dev)
        w, g = float(want.item()), float(got.item())
        assert g == pytest.approx(w, rel=1e-5), (call, nx, ny, w, g)
        # and both agree with a float64 evaluation on the host
        d = a[8:8 + nx, 32:32 + ny].double().cpu().numpy() - b[8:8 + nx, 32:32 + ny].double().cpu().numpy()
        ref = float((d ** 2).sum()) if norm == "l2sq" else float(np.abs(d).max())
        assert g == pytest.approx(ref, rel=1e-5)


def test_static_scratch_entry_point_rejects_oversized_tiles():
    lib = N.cuda()
    dev = torch.device("cuda:0")
    rows = N.NORM_STATIC_MAX_ROWS + 8
    tile = N.make_tile(rows, 4, 0, 0, rows, 4, 0, 0, 4)
    a = torch.zeros((rows, 4), dtype=torch.float32, device=dev)
    out = torch.zeros(1, dtype=torch.float32, device=dev)
    rc = lib.h2d_cuda_norm_diff_static(a.data_ptr(), a.data_ptr(), C.byref(tile), N.NORM["l2sq"],
                                       out.data_ptr(), int(torch.cuda.current_stream(dev).cuda_stream))
    assert rc == N.H2D_EINVAL
    assert "at most" in N.last_error(lib)
