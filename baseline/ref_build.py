"""Stage and build the UNMODIFIED reference programs under ``baseline/_ref/``.

The reference (patschris/Heat2D) is five loose source files with no build system
and no Python packaging: ``pip install /root/reference`` fails with "Neither
'setup.py' nor 'pyproject.toml' found" (recorded in DESIGN.md).  Its only "API"
is the ``#define`` block at the top of each file (``grad1612_cuda_heat.cu:6-13``,
``grad1612_mpi_heat.c:5-21``, ``mpi_heat2Dn.c:29-39``) plus the compile lines of
``readme.md:10-19``.  So "running the reference through its own public API" means:

1. copy the sources byte-for-byte into ``baseline/_ref/src/``  (``stage()``),
2. for a requested configuration, write a copy whose ``#define NAME value`` lines
   carry the requested values — the macros have no ``#ifndef`` guards, so ``-D``
   cannot override them — and nothing else changes (``configure()``),
3. compile it (``build_cuda()`` / ``build_mpi()``).

``baseline/_ref`` is git-ignored but ships to the GPU box; ``/root/reference``
does not exist there, which is why staging happens here first.
"""
from __future__ import annotations

import hashlib
import os
import re
import shutil
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF = HERE / "_ref"
SRC = REF / "src"
BUILD = REF / "build"
BIN = REF / "bin"
UPSTREAM = Path(os.environ.get("HEAT2D_REFERENCE", "/root/reference"))
FILES = ["grad1612_cuda_heat.cu", "grad1612_mpi_heat.c", "grad1612_hybrid_heat.c", "mpi_heat2Dn.c",
         "readme.md"]


def stage(force: bool = False) -> Path:
    """Copy the reference sources verbatim (needs /root/reference; run on the dev box)."""
    SRC.mkdir(parents=True, exist_ok=True)
    for f in FILES:
        dst = SRC / f
        if dst.exists() and not force:
            continue
        src = UPSTREAM / f
        if not src.exists():
            raise FileNotFoundError(f"{src} not found and {dst} not staged")
        shutil.copyfile(src, dst)
    sums = {f: hashlib.sha256((SRC / f).read_bytes()).hexdigest() for f in FILES}
    (REF / "SHA256SUMS").write_text("".join(f"{h}  {f}\n" for f, h in sums.items()))
    return SRC


def staged() -> bool:
    return all((SRC / f).exists() for f in FILES)


def configure(fname: str, defines: dict, tag: str) -> Path:
    """Copy of `fname` with the given ``#define`` values; everything else untouched."""
    text = (SRC / fname).read_text()
    for name, val in defines.items():
        pat = re.compile(r"^(#define[ \t]+%s[ \t]+)(\S+)" % re.escape(name), re.M)
        if not pat.search(text):
            raise KeyError(f"{fname} has no '#define {name}'")
        text = pat.sub(lambda m: m.group(1) + str(val), text, count=1)
    out_dir = BUILD / tag
    out_dir.mkdir(parents=True, exist_ok=True)
    out = out_dir / fname
    out.write_text(text)
    return out


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found")


def build_cuda(nx: int, ny: int, steps: int, readme_build: bool = False) -> Path:
    """grad1612_cuda_heat.cu for one problem size.  Default: native sm_100a SASS
    (fair to the reference); ``readme_build`` = the bare ``nvcc file.cu`` of
    readme.md:18 (sm_52 PTX, JIT-compiled by the driver)."""
    if not staged():
        stage()
    tag = f"cuda_{nx}x{ny}_{steps}" + ("_readme" if readme_build else "")
    BIN.mkdir(parents=True, exist_ok=True)
    exe = BIN / tag
    if exe.exists():
        return exe
    src = configure("grad1612_cuda_heat.cu", {"NXPROB": nx, "NYPROB": ny, "STEPS": steps}, tag)
    cmd = [_nvcc(), str(src), "-o", str(exe)]
    if not readme_build:
        cmd[1:1] = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("reference build failed:\n" + r.stdout + r.stderr)
    return exe


def run_cuda(exe: Path, gpu: int = 0, timeout: float = 3600.0) -> float:
    """Run the reference binary; return the seconds it prints
    (``Elapsed time: %e sec``, grad1612_cuda_heat.cu:89)."""
    env = dict(os.environ, CUDA_VISIBLE_DEVICES=str(gpu))
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=timeout, env=env)
    if r.returncode != 0:
        raise RuntimeError(f"{exe.name} exited {r.returncode}:\n{r.stdout}{r.stderr}")
    m = re.search(r"Elapsed time:\s*([0-9.eE+-]+)\s*sec", r.stdout)
    if not m:
        raise RuntimeError(f"no 'Elapsed time' line in:\n{r.stdout}")
    return float(m.group(1))


def main(argv=None):
    import argparse
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--stage", action="store_true")
    ap.add_argument("--cuda", nargs=3, type=int, metavar=("NX", "NY", "STEPS"), action="append", default=[])
    a = ap.parse_args(argv)
    if a.stage:
        print(stage(force=True))
    for nx, ny, st in a.cuda:
        print(build_cuda(nx, ny, st))


if __name__ == "__main__":
    main()
