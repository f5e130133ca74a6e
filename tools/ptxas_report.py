#!/usr/bin/env python
"""Summarise the ptxas -v logs of the last build (registers, spills, stack, smem per kernel)
into profiles/ptxas_summary.txt.  The build writes its logs under heat2d_b200/_lib/obj/
(untracked); this tool is the explicit step that turns them into a tracked record.

    python tools/ptxas_report.py [--all]      # default: the K = 8 kernels and the K = 4 strict ones
"""
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
OBJ = ROOT / "heat2d_b200" / "_lib" / "obj"


def demangle(names):
    try:
        out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
        return dict(zip(names, out))
    except OSError:
        return {n: n for n in names}


def main():
    want_all = "--all" in sys.argv
    rows = []
    for log in sorted(OBJ.glob("*.ptxas.log")):
        text = log.read_text().splitlines()
        for i, l in enumerate(text):
            m = re.search(r"Compiling entry function '(\S+)'", l)
            if not m:
                continue
            blob = " ".join(text[i:i + 4])
            regs = re.search(r"Used (\d+) registers", blob)
            spill = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", blob)
            smem = re.search(r"(\d+) bytes smem", blob)
            rows.append((log.stem.replace(".ptxas", ""), m.group(1), int(regs.group(1)) if regs else -1,
                         tuple(int(x) for x in spill.groups()) if spill else (0, 0, 0),
                         int(smem.group(1)) if smem else 0))
    names = demangle([r[1] for r in rows])
    lines = ["# ptxas -v summary (sm_100a): object | kernel | registers | stack / spill-store / spill-load bytes | static smem"]
    for obj, mangled, regs, sp, smem in rows:
        name = names[mangled]
        name = re.sub(r"\(CUtensorMap_st.*", "", name).replace("h2d::", "").replace("void ", "")
        # K = 8 everywhere, plus K = 4 of the strict modes (bench.py's headline runs there)
        if not want_all and "tk_" in name and not re.search(r"<8,|<4, [12],", name):
            continue
        lines.append(f"{obj:16s} {name:58s} {regs:4d}  {sp[0]:4d}/{sp[1]:4d}/{sp[2]:4d}  {smem}")
    out = ROOT / "profiles" / "ptxas_summary.txt"
    out.write_text("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main()
