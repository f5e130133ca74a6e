#!/usr/bin/env python
"""Per-loop instruction mix of one kernel's SASS (cuobjdump -sass output on stdin or a file):
every backward branch closes a loop; nested loops are reported separately.  Used to check
that spill code (LDL/STL) and register moves stay out of the hot row loops.

    cuobjdump -sass -fun <mangled> lib.so | python tools/sass_loops.py [min_instructions]
"""
import collections
import re
import sys


def main():
    import signal
    signal.signal(signal.SIGPIPE, signal.SIG_DFL)
    src = sys.stdin if len(sys.argv) < 2 or not sys.argv[-1].endswith(".sass") else open(sys.argv[-1])
    min_n = next((int(a) for a in sys.argv[1:] if a.isdigit()), 200)
    rows = []
    for l in src:
        m = re.match(r"\s*/\*([0-9a-f]{4,6})\*/\s+(.*?);", l)
        if m:
            rows.append((int(m.group(1), 16), m.group(2).strip()))
    addr_index = {a: i for i, (a, _) in enumerate(rows)}
    for i, (a, t) in enumerate(rows):
        m = re.search(r"\bBRA(?:\.\S+)?\s+(?:!?U?P\d+,\s*)?(0x[0-9a-f]+)", t)
        if not m:
            continue
        tgt = int(m.group(1), 16)
        if tgt >= a or tgt not in addr_index:
            continue
        body = rows[addr_index[tgt]:i + 1]
        if len(body) < min_n:
            continue
        cnt = collections.Counter()
        for _, ins in body:
            op = re.sub(r"^@!?U?P\d+\s+", "", ins).split()[0].split(".")[0]
            cnt[op] += 1
        keys = ["FFMA2", "FADD2", "FMUL2", "FFMA", "FADD", "FMUL", "DFMA", "DADD", "DMUL", "F2F", "SHFL", "FSEL",
                "SEL", "MOV", "IMAD", "SHF", "LDS", "STG", "LDL", "STL", "ISETP", "BRA", "VIMNMX", "IMNMX"]
        mix = " ".join(f"{k}={cnt[k]}" for k in keys if cnt[k])
        print(f"loop {tgt:#x}..{a:#x}: {len(body)} instr | {mix}")


if __name__ == "__main__":
    main()
