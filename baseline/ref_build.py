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


SHIM_SRC = HERE / "mpishim"
SHIM = REF / "shim"


def _gcc() -> str:
    for cand in ("/usr/bin/gcc", os.environ.get("CC"), shutil.which("gcc")):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("gcc not found")


def build_shim(force: bool = False) -> Path:
    """Build the single-node MPI shim (baseline/mpishim) into baseline/_ref/shim/."""
    SHIM.mkdir(parents=True, exist_ok=True)
    lib, exe = SHIM / "libmpi.a", SHIM / "mpiexec"
    srcs = [SHIM_SRC / "mpishim.c", SHIM_SRC / "mpiexec.c", SHIM_SRC / "mpi.h"]
    if not force and lib.exists() and exe.exists() and \
            min(lib.stat().st_mtime, exe.stat().st_mtime) >= max(f.stat().st_mtime for f in srcs):
        return SHIM
    cc = _gcc()
    obj = SHIM / "mpishim.o"
    for cmd in ([cc, "-O2", "-Wall", f"-I{SHIM_SRC}", "-c", str(SHIM_SRC / "mpishim.c"), "-o", str(obj)],
                ["ar", "rcs", str(lib), str(obj)],
                [cc, "-O2", "-Wall", str(SHIM_SRC / "mpiexec.c"), "-o", str(exe)]):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("shim build failed:\n" + r.stdout + r.stderr)
    return SHIM


MPI_PROGRAMS = {"mpi_heat2Dn": "mpi_heat2Dn.c", "grad1612_mpi": "grad1612_mpi_heat.c",
                "grad1612_hybrid": "grad1612_hybrid_heat.c"}


def build_mpi(program: str, defines: dict, opt: str = "-O3") -> Path:
    """One of the three MPI reference programs on the shim.  `opt` replaces the
    README's ``-g`` (readme.md:10,14); the mpiP build of the report used -O3."""
    if not staged():
        stage()
    build_shim()
    fname = MPI_PROGRAMS[program]
    tag = program + "_" + "_".join(f"{k}{v}" for k, v in sorted(defines.items())) + "_" + opt.strip("-")
    BIN.mkdir(parents=True, exist_ok=True)
    exe = BIN / tag
    if exe.exists():
        return exe
    src = configure(fname, defines, tag)
    cmd = [_gcc(), opt, "-Wall", "-Wno-unused-result", "-Wno-unused-variable", "-Wno-implicit-int",
           "-Wno-unused-but-set-variable", "-Wno-maybe-uninitialized", "-Wno-deprecated-non-prototype",
           f"-I{SHIM_SRC}", str(src), str(SHIM / "libmpi.a"), "-o", str(exe)]
    if program == "grad1612_hybrid":
        cmd.insert(2, "-fopenmp")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("reference MPI build failed:\n" + r.stdout + r.stderr)
    return exe


def run_mpi(exe: Path, nproc: int, cwd, timeout: float = 3600.0, env_extra=None):
    """mpiexec -n nproc exe, in `cwd` (the programs write initial.dat / final.dat there).
    Returns (stdout, elapsed seconds printed by the program or None)."""
    build_shim()
    env = dict(os.environ)
    if env_extra:
        env.update(env_extra)
    r = subprocess.run([str(SHIM / "mpiexec"), "-n", str(nproc), str(exe)], capture_output=True, text=True,
                       timeout=timeout, cwd=str(cwd), env=env)
    if r.returncode != 0:
        raise RuntimeError(f"{exe.name} on {nproc} ranks exited {r.returncode}:\n{r.stdout[-2000:]}{r.stderr[-2000:]}")
    m = re.search(r"Elapsed time:\s*([0-9.eE+-]+)\s*sec", r.stdout)
    return r.stdout, (float(m.group(1)) if m else None)


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
