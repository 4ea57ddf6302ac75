#!/usr/bin/env python
"""SASS opcode tally + register / spill table of every kernel of the built library (evidence for profiles/):
     python tools/sass_summary.py > profiles/<round>_sass_summary.txt
Memory opcodes are listed with their width suffix (LDG.E.128 = 128-bit global load ...); UBLKCP = cp.async.bulk (TMA bulk
copy), SYNCS = mbarrier operations; LDL / STL = local-memory (spill) accesses."""
import collections
import glob
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return {n: re.sub(r"\(.*", "", d).replace("fbn::", "").replace("(anonymous namespace)::", "") for n, d in zip(names, out)}


def main():
    print("== registers / spills per kernel (nvcc -Xptxas -v, build/csrc/*.ptxas.log)")
    rows = []
    for f in sorted(glob.glob(os.path.join(ROOT, "build", "csrc", "*.ptxas.log"))):
        name = None
        spill = ""
        for line in open(f):
            m = re.search(r"Compiling entry function '([^']+)'", line)
            if m:
                name = m.group(1)
            if "bytes stack frame" in line:
                spill = line.strip()
            m = re.search(r"Used (\d+) registers", line)
            if m and name:
                rows.append((name, int(m.group(1)), spill))
                name = None
    dm = demangle([r[0] for r in rows])
    for n, regs, spill in rows:
        print("  %-44s %3d registers   %s" % (dm[n], regs, spill))
    print()
    print("== SASS opcode tally per kernel (cuobjdump -sass fastbn_b200/libfastbn_b200.so)")
    sass = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, "fastbn_b200", "libfastbn_b200.so")], capture_output=True, text=True).stdout
    cnt = collections.defaultdict(collections.Counter)
    name = None
    for line in sass.split("\n"):
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and name:
            op = m.group(1)
            cnt[name]["TOTAL"] += 1
            if re.match(r"(LDG|STG|LDL|STL|LDS|STS|UBLKCP|UTMA|SYNCS|ATOM|RED)", op):
                cnt[name][op] += 1
            elif re.match(r"(DFMA|DMUL|DADD|SHFL|BAR)", op):
                cnt[name][op.split(".")[0]] += 1
    dm = demangle(list(cnt))
    for n in sorted(cnt, key=lambda x: dm[x]):
        c = cnt[n]
        print("  %s   (%d instructions)" % (dm[n], c["TOTAL"]))
        keys = sorted(k for k in c if k != "TOTAL")
        print("      " + "  ".join("%s %d" % (k, c[k]) for k in keys))


if __name__ == "__main__":
    main()
