"""The C-ABI example builds against the in-tree library (and runs on a GPU box)."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(tmp_path):
    from heat2d_b200 import build as B
    B.build_cuda()
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    exe = tmp_path / "c_api_demo"
    lib = os.path.join(ROOT, "heat2d_b200", "_lib")
    r = subprocess.run([nvcc, "-O2", "-gencode", "arch=compute_100a,code=sm_100a", "-I" + os.path.join(ROOT, "include"),
                        os.path.join(ROOT, "examples", "c_api_demo.cu"), "-L" + lib, "-lheat2d_cuda",
                        "-Xlinker", "-rpath", "-Xlinker", lib, "-o", str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_c_example_compiles_and_links(tmp_path):
    assert _build(tmp_path).exists()


@pytest.mark.gpu
def test_c_example_runs(tmp_path):
    exe = _build(tmp_path)
    r = subprocess.run([str(exe), "640", "512", "100"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Problem size: 640x512" in r.stdout and "Elapsed time:" in r.stdout
    m = re.search(r"u\(centre\) = ([0-9.e+-]+)", r.stdout)
    import numpy as np
    from heat2d_b200 import _native as N
    from heat2d_b200.ops import kernels as K
    ref = K.cpu_initial_grid(640, 512, "exact")
    K.cpu_run(ref, 100, N.make_phys())
    assert float(m.group(1)) == pytest.approx(float(ref[320, 256]), rel=1e-5)
