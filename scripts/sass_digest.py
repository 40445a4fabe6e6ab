#!/usr/bin/env python
"""profiles/sass_digest.txt: per kernel of canvas_b200/libcanvas_b200.so, the counts of the SASS opcodes that show
whether the hot path is Blackwell-native and whether anything regressed (bulk copies UBLKCP, mbarrier ops SYNCS, packed
FP32 FFMA2/FMUL2/FADD2, 128-bit shared loads and global stores, MUFU.RCP, local-memory spills LDL/STL), plus registers
per thread from the cubin's resource usage.  Regenerate after every kernel change:

    python scripts/sass_digest.py            # writes profiles/sass_digest.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "canvas_b200", "libcanvas_b200.so")
COLUMNS = ["UBLKCP", "SYNCS", "FFMA2", "FMUL2", "FADD2", "FFMA", "MUFU.RCP", "LDS.128", "LDS.U16", "LDG", "STG.128", "STG", "SHFL", "PRMT", "LDL", "STL", "total"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    sass = subprocess.run(["cuobjdump", "-sass", "-arch", "sm_100a", LIB], capture_output=True, text=True, check=True).stdout
    usage = subprocess.run(["cuobjdump", "-res-usage", "-arch", "sm_100a", LIB], capture_output=True, text=True).stdout
    regs = {}
    current = None
    for line in usage.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            current = m.group(1)
        m = re.search(r"REG:(\d+).*?SHARED:(\d+)", line)
        if m and current:
            regs[current] = (int(m.group(1)), int(m.group(2)))

    counts, order = {}, []
    current = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            current = m.group(1)
            counts[current] = collections.Counter()
            order.append(current)
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and current:
            op = m.group(1)
            c = counts[current]
            c["total"] += 1
            for key in COLUMNS[:-1]:
                if key in ("LDG", "STG", "FFMA"):
                    if op.split(".")[0] == key:
                        c[key] += 1
                elif key == "STG.128":
                    if op.startswith("STG") and ".128" in op:
                        c[key] += 1
                elif key == "LDS.128":
                    if op.startswith("LDS") and ".128" in op:
                        c[key] += 1
                elif key == "LDS.U16":
                    if op.startswith("LDS") and ".U16" in op:
                        c[key] += 1
                elif op.startswith(key):
                    c[key] += 1

    names = demangle(order)
    lines = ["# SASS opcode digest of canvas_b200/libcanvas_b200.so (cuobjdump -sass -arch sm_100a), by scripts/sass_digest.py",
             "# columns: registers/thread, " + ", ".join(COLUMNS),
             ""]
    totals = collections.Counter()
    for fn in sorted(order, key=lambda f: names[f]):
        c = counts[fn]
        totals.update(c)
        short = re.sub(r"\(.*", "", names[fn]).replace("void ", "")
        lines.append("%-58s regs %3s  " % (short, regs.get(fn, ("?",))[0]) + " ".join("%s=%d" % (k, c[k]) for k in COLUMNS if c[k]))
    lines.append("")
    lines.append("ALL KERNELS (%d): " % len(order) + " ".join("%s=%d" % (k, totals[k]) for k in COLUMNS))
    spills = [names[f] for f in order if counts[f]["LDL"] or counts[f]["STL"]]
    lines.append("kernels with local-memory traffic (spills or indexed local arrays): %d" % len(spills))
    for s in spills:
        lines.append("  " + re.sub(r"\(.*", "", s))
    out = os.path.join(ROOT, "profiles", "sass_digest.txt")
    with open(out, "w") as f:
        f.write("\n".join(lines) + "\n")
    print("\n".join(lines[-(len(spills) + 3):]))


if __name__ == "__main__":
    sys.exit(main())
