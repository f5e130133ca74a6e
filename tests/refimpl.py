"""Independent numpy restatement of the reference arithmetic, used to check the
C++ oracle (which in turn checks the GPU kernels)."""
import numpy as np


def np_init(nx, ny, mode="exact"):
    x = np.arange(nx, dtype=np.int64)[:, None]
    y = np.arange(ny, dtype=np.int64)[None, :]
    if mode == "exact":
        return (x * (nx - x - 1) * y * (ny - y - 1)).astype(np.float32)
    with np.errstate(over="ignore"):
        xu = x.astype(np.uint32)
        yu = y.astype(np.uint32)
        p = (xu * np.uint32(1)) * (np.uint32(nx) - xu - np.uint32(1))
        p = p.astype(np.uint32) * yu
        p = p.astype(np.uint32) * (np.uint32(ny) - yu - np.uint32(1))
        p = p.astype(np.uint32)
    if mode == "ref-uint32":
        return p.astype(np.float32)
    return p.view(np.int32).astype(np.float32)


def np_step(u, cx=0.1, cy=0.1, arith="strict", boundary="fixed"):
    nx, ny = u.shape
    p = np.pad(u.astype(np.float32), 1)
    c = p[1:-1, 1:-1]
    xm, xp = p[:-2, 1:-1], p[2:, 1:-1]
    ym, yp = p[1:-1, :-2], p[1:-1, 2:]
    ew = (xp + xm).astype(np.float32)
    ns = (yp + ym).astype(np.float32)
    if arith == "fp32":
        # no fused multiply-add in numpy: this is the float64 value of the same
        # expression, to be compared with a tolerance
        a0 = np.float32(1.0 - 2.0 * cx - 2.0 * cy)
        new = (np.float64(a0) * c + np.float64(np.float32(cy)) * ns + np.float64(np.float32(cx)) * ew)
        new = new.astype(np.float32)
    else:
        c64 = c.astype(np.float64)
        t1 = ew.astype(np.float64) - 2.0 * c64
        t2 = ns.astype(np.float64) - 2.0 * c64
        new = ((c64 + cx * t1) + cy * t2).astype(np.float32)
    out = u.copy()
    if boundary == "fixed":
        out[1:-1, 1:-1] = new[1:-1, 1:-1]
    else:
        out[:, :] = new
    return out


def np_run(u, steps, **kw):
    for _ in range(steps):
        u = np_step(u, **kw)
    return u
