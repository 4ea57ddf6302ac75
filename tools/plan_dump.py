#!/usr/bin/env python
"""Developer tool: structure of the device plan of a workload (no GPU needed).  python tools/plan_dump.py <workload> [team] [level] [verbose]"""
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import emu  # noqa: E402

so = os.path.join(ROOT, "build", "tools", "libplan_dump.so")
os.makedirs(os.path.dirname(so), exist_ok=True)
subprocess.check_call(["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "fastbn_b200", "csrc"),
                       os.path.join(ROOT, "tools", "plan_dump.cpp"), os.path.join(ROOT, "fastbn_b200", "csrc", "plan.cpp"), os.path.join(ROOT, "fastbn_b200", "csrc", "resident.cpp"), "-shared", "-fPIC", "-o", so])
w = sys.argv[1]
team = int(sys.argv[2]) if len(sys.argv) > 2 else 1
level = int(sys.argv[3]) if len(sys.argv) > 3 else 0
verbose = int(sys.argv[4]) if len(sys.argv) > 4 else 1
d, keep = emu.tree_desc(bench.load_tree(w))
sys.stdout.flush()
L = C.CDLL(so)
if team == 0:
    L.gather_stats(C.byref(d))
elif team < 0:
    L.resident_dump(C.byref(d), -team, verbose)  # negative team: the resident-tile schedule for -team warps
else:
    L.plan_dump(C.byref(d), team, level, verbose)
