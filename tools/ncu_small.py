"""Launch every small kernel once (init, norm, pack/unpack, resident, finish) for an ncu capture."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from heat2d_b200 import HeatConfig, _native as N
from heat2d_b200.models.heat import Heat2DSolver, TileField
from heat2d_b200.ops import kernels as K
from heat2d_b200.parallel.decomp import Decomposition

dev = "cuda:0"
n = 8192
d = Decomposition(n, n, 1, 1)
f = TileField(d.tile(0), n, n, 8, dev)
for _ in range(2):
    K.init_field(f.buf[0], f.tile, "exact", True)
    K.init_field(f.buf[1], f.tile, "ref-uint32", True)
    v = K.norm_diff(f.buf[0], f.buf[1], f.tile, "l2sq")
    buf = torch.empty(n, dtype=torch.float32, device=dev)
    K.pack_edge(f.buf[0], f.tile, "W", 1, buf)
    K.unpack_ghost(f.buf[0], f.tile, "E", 1, buf)
torch.cuda.synchronize()
# convergence path: fused norm + finish kernel
s = Heat2DSolver(HeatConfig(nx=4096, ny=4096, steps=21, k=8, conv=True, interval=20, eps=0.0), dev).initialize()
s.run()
# resident cluster kernel
r = Heat2DSolver(HeatConfig(nx=640, ny=512, steps=1000, backend="resident"), dev).initialize()
res = r.run()
print("resident 640x512 x1000:", res.elapsed_s, "s", res.glups, "GLUP/s")
r2 = Heat2DSolver(HeatConfig(nx=80, ny=64, steps=1000, backend="resident"), dev).initialize()
res2 = r2.run()
print("resident 80x64 x1000:", res2.elapsed_s, "s", res2.glups, "GLUP/s", res2.elapsed_s / 1000 * 1e6, "us/step")
for nx, ny in [(80, 64), (160, 128), (320, 256), (640, 512), (1280, 1024), (2560, 2048)]:
    t = Heat2DSolver(HeatConfig(nx=nx, ny=ny, steps=1000, k=8), dev).initialize()
    t.run(64)
    rr = t.run(1000)
    print(f"fused k=8 {nx}x{ny} x1000: {rr.elapsed_s:.6f} s  {rr.glups:.2f} GLUP/s")
torch.cuda.synchronize()
