"""Build the two native libraries in-tree.

    python -m heat2d_b200.build [--force] [--verbose]

* ``_lib/libheat2d_cpu.so``  — oracle + dump writers, ``g++ -fopenmp -ffp-contract=off``
* ``_lib/libheat2d_cuda.so`` — every CUDA kernel and the peer-memory runtime,
  ``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo`` (sm_100a only; nvcc
  cross-compiles without a GPU).

The ``.so`` files are git-ignored but travel to the GPU box with the snapshot.
The reference has no build system at all (three command lines in readme.md:10-19).
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import contextlib
import fcntl
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
INC = ROOT / "include"
LIB = Path(__file__).resolve().parent / "_lib"
OBJ = LIB / "obj"

CPU_SOURCES = ["oracle.cpp", "dump.cpp"]
CUDA_SOURCES = ["basic_kernels.cu", "stencil_tk.cu", "stencil_flow.cu", "stencil_team.cu", "resident.cu", "peer.cu"]
# tk_inst.cu is compiled once per (arithmetic, lane layout, extras) combination:
# (object name, ARITH, LAY, XTR).  ARITH 0 fp32 / 1 strict / 2 strict-fma (heat2d.h);
# LAY 0 v4 / 1 ab2 / 2 ab4 (csrc/tk_task.cuh); strict arithmetic exists in the v4 layout only.
TK_INSTANCES = [(f"tk_a{a}_l{l}_x{x}", a, l, x)
                for a, ls in ((0, (0, 1, 2)), (1, (0, 3)), (2, (0, 3))) for l in ls for x in (0, 1)]
# defaults baked into the library (both can be overridden at run time, see stencil_tk.cu)
TK_DEFAULTS = ["-DH2D_DEFAULT_FP32_LAYOUT=0", "-DH2D_DEFAULT_MIN_WAVES=2", "-DH2D_DEFAULT_STRICT_SHALLOW_LAYOUT=3"]

NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v",
]
CXX_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-fopenmp", "-ffp-contract=off", "-Wall"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found")


def _newer(target: Path, deps) -> bool:
    """True if `target` exists and no source is newer.  An installed copy of the package has
    the libraries but not the sources (csrc/, include/): the libraries are then used as is."""
    if not target.exists():
        return False
    t = target.stat().st_mtime
    return all(not Path(d).exists() or Path(d).stat().st_mtime <= t for d in deps)


def _headers():
    return list(CSRC.glob("*.h")) + list(CSRC.glob("*.cuh")) + list(INC.glob("*.h"))


@contextlib.contextmanager
def _build_lock():
    """One builder at a time.  Under torchrun every rank imports the package at once on a
    fresh checkout; without the lock they would write the same objects concurrently and a
    rank could dlopen a half-written library.  Outputs are also written to a temporary
    name and renamed into place, so a reader never sees a partial file."""
    LIB.mkdir(parents=True, exist_ok=True)
    with open(LIB / ".build.lock", "w") as fh:
        fcntl.flock(fh, fcntl.LOCK_EX)
        try:
            yield
        finally:
            fcntl.flock(fh, fcntl.LOCK_UN)


def _run(cmd, verbose, log_path=None):
    res = subprocess.run(cmd, capture_output=True, text=True)
    if log_path is not None:
        # ptxas -v output (registers / spills); `python tools/ptxas_report.py` copies a
        # summary under profiles/ — the build itself writes nothing into tracked files
        text = "\n".join(l for l in (res.stdout + res.stderr).splitlines() if "Compile time" not in l)
        Path(log_path).write_text(text + "\n")
    if res.returncode != 0:
        raise RuntimeError("build failed: %s\n%s\n%s" % (" ".join(map(str, cmd)), res.stdout, res.stderr))
    if verbose:
        sys.stderr.write(res.stdout + res.stderr)


def build_cpu(force=False, verbose=False) -> Path:
    LIB.mkdir(parents=True, exist_ok=True)
    out = LIB / "libheat2d_cpu.so"
    srcs = [CSRC / s for s in CPU_SOURCES]
    if not force and _newer(out, srcs + _headers()):
        return out
    with _build_lock():
        if not force and _newer(out, srcs + _headers()):   # another process built it meanwhile
            return out
        return _build_cpu_locked(out, srcs, verbose)


def _build_cpu_locked(out: Path, srcs, verbose) -> Path:
    # Some images export CXX pointing at a g++ without libgomp; try the system
    # compiler first and drop -fopenmp only as a last resort (the oracle then
    # runs serially, results are identical).
    cands = [c for c in ("/usr/bin/g++", os.environ.get("CXX"), shutil.which("g++")) if c]
    errors = []
    for flags in (CXX_FLAGS, [f for f in CXX_FLAGS if f != "-fopenmp"]):
        for cxx in dict.fromkeys(cands):
            tmp = out.with_suffix(f".so.tmp{os.getpid()}")
            cmd = [cxx, *flags, "-shared", f"-I{INC}", f"-I{CSRC}", *map(str, srcs), "-o", str(tmp)]
            try:
                _run(cmd, verbose)
                os.replace(tmp, out)
                return out
            except RuntimeError as e:  # try the next compiler
                errors.append(str(e))
    raise RuntimeError("no working C++ compiler:\n" + "\n".join(errors))


def build_cuda(force=False, verbose=False) -> Path:
    # Experiment hook: H2D_VARIANT=name H2D_NVCC_EXTRA="-DFOO=1" builds
    # libheat2d_cuda_name.so next to the default library (tools/ only).
    variant = os.environ.get("H2D_VARIANT", "")
    extra = os.environ.get("H2D_NVCC_EXTRA", "").split()
    global OBJ
    LIB.mkdir(parents=True, exist_ok=True)
    obj_dir = LIB / ("obj_" + variant if variant else "obj")
    OBJ = obj_dir
    OBJ.mkdir(parents=True, exist_ok=True)
    out = LIB / ("libheat2d_cuda_%s.so" % variant if variant else "libheat2d_cuda.so")
    srcs = [CSRC / s for s in CUDA_SOURCES]
    hdrs = _headers()
    if not force and _newer(out, srcs + [CSRC / "tk_inst.cu"] + hdrs):
        return out
    nvcc = _nvcc()

    def compile_one(job) -> Path:
        src, name, defs = job
        obj = OBJ / (name + ".o")
        if not force and _newer(obj, [src] + hdrs):
            return obj
        tmp = OBJ / (name + f".o.tmp{os.getpid()}")
        cmd = [nvcc, *NVCC_FLAGS, *TK_DEFAULTS, *defs, *extra, f"-I{INC}", f"-I{CSRC}", "-c", str(src), "-o", str(tmp)]
        _run(cmd, verbose, log_path=OBJ / (name + ".ptxas.log"))
        os.replace(tmp, obj)
        return obj

    # the instantiation units first: they are the long poles
    jobs = [(CSRC / "tk_inst.cu", name, [f"-DH2D_INST_ARITH={a}", f"-DH2D_INST_LAY={l}", f"-DH2D_INST_XTR={x}"])
            for name, a, l, x in TK_INSTANCES]
    jobs += [(src, src.stem, []) for src in srcs]
    srcs = srcs + [CSRC / "tk_inst.cu"]
    with _build_lock():
        if not force and _newer(out, srcs + hdrs):   # another rank built it while we waited
            return out
        with cf.ThreadPoolExecutor(max_workers=max(1, min(os.cpu_count() or 4, len(jobs)))) as ex:
            objs = list(ex.map(compile_one, jobs))
        tmp = out.with_suffix(f".so.tmp{os.getpid()}")
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", *map(str, objs), "-o", str(tmp)]
        _run(cmd, verbose)
        os.replace(tmp, out)
    return out


def build_all(force=False, verbose=False):
    return build_cpu(force, verbose), build_cuda(force, verbose)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--cpu-only", action="store_true")
    args = ap.parse_args(argv)
    cpu = build_cpu(args.force, args.verbose)
    print(cpu)
    if not args.cpu_only:
        print(build_cuda(args.force, args.verbose))


if __name__ == "__main__":
    main()
