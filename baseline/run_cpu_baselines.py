"""Time the reference's three CPU programs (unmodified sources, single-node MPI shim)
on this machine's cores and write baseline/_ref/cpu_baselines.json.

    python baseline/run_cpu_baselines.py [--nx 2560 --ny 2048 --steps 1000]

Numbers are "reference source on shim-MPI, gcc -O3" — not OpenMPI, not the
cluster of the report.
"""
import argparse
import json
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ref_build as rb  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nx", type=int, default=2560)
    ap.add_argument("--ny", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--out", default=os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "profiles",
                                                  "cpu_reference_on_shim.json"))
    a = ap.parse_args()
    ncpu = os.cpu_count() or 1
    out = {"grid": [a.nx, a.ny], "steps": a.steps, "cpus": ncpu, "label": "reference source on shim-MPI, gcc -O3",
           "when": time.strftime("%Y-%m-%dT%H:%M:%S"), "runs": []}

    def rec(program, ranks, grid, secs, **kw):
        glups = a.nx * a.ny * a.steps / secs / 1e9 if secs else None
        out["runs"].append(dict(program=program, ranks=ranks, process_grid=grid, seconds=secs, glups=glups, **kw))
        print(out["runs"][-1], flush=True)

    # grad1612_mpi_heat.c, no convergence (Report.pdf Table 1)
    for gx, gy in [(1, 1), (2, 2), (2, 4)]:
        if gx * gy > ncpu:
            continue
        exe = rb.build_mpi("grad1612_mpi", {"NXPROB": a.nx, "NYPROB": a.ny, "STEPS": a.steps, "GRIDX": gx, "GRIDY": gy})
        with tempfile.TemporaryDirectory() as d:
            _, secs = rb.run_mpi(exe, gx * gy, d)
        rec("grad1612_mpi_heat.c", gx * gy, [gx, gy], secs, convergence=False)
    # with the convergence check (Table 4): as shipped the predicate fires every step iff xcell % 20 == 0
    for gx, gy in [(1, 1), (2, 2)]:
        if gx * gy > ncpu:
            continue
        exe = rb.build_mpi("grad1612_mpi", {"NXPROB": a.nx, "NYPROB": a.ny, "STEPS": a.steps, "GRIDX": gx, "GRIDY": gy,
                                            "CONVERGENCE": 1})
        with tempfile.TemporaryDirectory() as d:
            _, secs = rb.run_mpi(exe, gx * gy, d)
        rec("grad1612_mpi_heat.c", gx * gy, [gx, gy], secs, convergence=True,
            note="shipped predicate: checks every step iff (NX/GRIDX) % 20 == 0")
    # hybrid, 1 rank x 4 threads (Table 7: the nested-parallel defect makes it slower than 1 MPI rank)
    exe = rb.build_mpi("grad1612_hybrid", {"NXPROB": a.nx, "NYPROB": a.ny, "STEPS": a.steps, "CONVERGENCE": 0})
    with tempfile.TemporaryDirectory() as d:
        _, secs = rb.run_mpi(exe, 1, d)
    rec("grad1612_hybrid_heat.c", 1, [1, 1], secs, threads=4, convergence=False)
    # the given program: 3 and 7 workers (+ idle master); its stack arrays limit the grid
    for workers in (3, 7):
        if workers + 1 > ncpu:
            continue
        exe = rb.build_mpi("mpi_heat2Dn", {"NXPROB": 640, "NYPROB": 512, "STEPS": a.steps})
        with tempfile.TemporaryDirectory() as d:
            _, secs = rb.run_mpi(exe, workers + 1, d, env_extra={})
        out["runs"].append(dict(program="mpi_heat2Dn.c", ranks=workers + 1, process_grid=[workers, 1], seconds=secs,
                                grid=[640, 512], glups=640 * 512 * a.steps / secs / 1e9 if secs else None))
        print(out["runs"][-1], flush=True)
    os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
    with open(a.out, "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", a.out)


if __name__ == "__main__":
    main()
