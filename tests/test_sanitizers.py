"""ASan / UBSan builds of the host-side native code ("race detection /
sanitizers"): the CPU oracle + dump writers, and the MPI shim the reference programs are
linked against.  The reference itself ships no sanitizer build; its ``inidat`` overflows a
signed int from 432^2 up (mpi_heat2Dn.c:247), which is why the shim test stays at 10x10.

CUDA kernels cannot be checked here (no GPU; compute-sanitizer needs one) — see
tools/sanitize_target.py for the racecheck/memcheck target.
"""
import os
import shutil
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
SAN = ["-fsanitize=address,undefined", "-fno-sanitize-recover=all", "-g", "-O1"]
ENV = dict(os.environ, ASAN_OPTIONS="detect_leaks=1:abort_on_error=0", UBSAN_OPTIONS="print_stacktrace=1")


def _cc(name):
    for cand in (f"/usr/bin/{name}", shutil.which(name)):
        if cand and Path(cand).exists():
            return cand
    pytest.skip(f"{name} not found")


def _sanitizers_work(tmp_path, cc, ext):
    src = tmp_path / f"probe.{ext}"
    src.write_text("int main(void){return 0;}\n")
    r = subprocess.run([cc, *SAN, str(src), "-o", str(tmp_path / "probe")], capture_output=True, text=True)
    return r.returncode == 0 and subprocess.run([str(tmp_path / "probe")], env=ENV).returncode == 0


def test_cpu_library_under_asan_ubsan(tmp_path):
    cxx = _cc("g++")
    if not _sanitizers_work(tmp_path, cxx, "cpp"):
        pytest.skip("sanitizer runtime not installed")
    exe = tmp_path / "sanitize_driver"
    cmd = [cxx, "-std=c++17", "-ffp-contract=off", *SAN, f"-I{ROOT / 'include'}", f"-I{ROOT / 'csrc'}",
           str(ROOT / "csrc" / "oracle.cpp"), str(ROOT / "csrc" / "dump.cpp"),
           str(ROOT / "tests" / "native" / "sanitize_driver.cpp"), "-o", str(exe)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    out = tmp_path / "out"
    out.mkdir()
    r = subprocess.run([str(exe), str(out)], capture_output=True, text=True, env=ENV, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "sanitize_driver: ok" in r.stdout


@pytest.mark.parametrize("program,nproc,defines", [
    ("grad1612_mpi_heat.c", 4, {"NXPROB": 10, "NYPROB": 10, "STEPS": 25, "GRIDX": 2, "GRIDY": 2}),
    ("grad1612_mpi_heat.c", 2, {"NXPROB": 12, "NYPROB": 10, "STEPS": 60, "GRIDX": 2, "GRIDY": 1,
                                "CONVERGENCE": 1}),
    ("mpi_heat2Dn.c", 4, {"NXPROB": 10, "NYPROB": 10, "STEPS": 25}),
])
def test_mpi_shim_under_asan_ubsan(tmp_path, program, nproc, defines):
    """The shim (instrumented) carrying an unmodified reference program (not instrumented):
    every pack/unpack of a derived datatype, persistent request and file view runs under
    ASan.  The result must equal the plain build's."""
    sys.path.insert(0, str(ROOT / "baseline"))
    import ref_build
    if not Path("/root/reference").exists() and not ref_build.staged():
        pytest.skip("reference sources not available")
    cc = _cc("gcc")
    if not _sanitizers_work(tmp_path, cc, "c"):
        pytest.skip("sanitizer runtime not installed")
    if not ref_build.staged():
        ref_build.stage()
    shim_src = ROOT / "baseline" / "mpishim"
    obj, launcher = tmp_path / "mpishim.o", tmp_path / "mpiexec"
    for cmd in ([cc, *SAN, "-Wall", f"-I{shim_src}", "-c", str(shim_src / "mpishim.c"), "-o", str(obj)],
                [cc, *SAN, "-Wall", str(shim_src / "mpiexec.c"), "-o", str(launcher)]):
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
    src = ref_build.configure(program, defines, "asan_" + "_".join(f"{k}{v}" for k, v in sorted(defines.items())))
    exe = tmp_path / "prog"
    cmd = [cc, "-O1", "-g", "-w", f"-I{shim_src}", "-c", str(src), "-o", str(tmp_path / "prog.o")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([cc, *SAN, str(tmp_path / "prog.o"), str(obj), "-o", str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr

    run_dir = tmp_path / "run"
    run_dir.mkdir()
    # Leak checking is off for this one: the reference never frees its persistent
    # requests (grad1612_mpi_heat.c:301-314) nor, on early exit, anything else.
    env = dict(ENV, ASAN_OPTIONS="detect_leaks=0")
    r = subprocess.run([str(launcher), "-n", str(nproc), str(exe)], capture_output=True, text=True,
                       cwd=str(run_dir), env=env, timeout=300)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "ERROR: AddressSanitizer" not in r.stderr and "runtime error" not in r.stderr, r.stderr[-3000:]
    assert "Elapsed time" in r.stdout

    # Same program on the plain shim: identical dump.  For the grad1612 program on several
    # ranks only initial.dat is comparable: its final dump omits MPI_File_set_view
    # (grad1612_mpi_heat.c:283-285), so every rank writes at offset 0 and the file
    # depends on who wrote last.
    name = {"grad1612_mpi_heat.c": "grad1612_mpi", "mpi_heat2Dn.c": "mpi_heat2Dn"}[program]
    plain = ref_build.build_mpi(name, defines, opt="-O1")
    plain_dir = tmp_path / "plain"
    plain_dir.mkdir()
    out_plain, _ = ref_build.run_mpi(plain, nproc, plain_dir, timeout=300)
    which = "final.dat" if name == "mpi_heat2Dn" else "initial.dat"
    assert (run_dir / which).read_bytes() == (plain_dir / which).read_bytes()
    if defines.get("CONVERGENCE"):
        pick = lambda text: [l for l in text.splitlines() if "Exiting after" in l]
        assert pick(r.stdout) == pick(out_plain) and pick(out_plain)
