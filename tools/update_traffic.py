#!/usr/bin/env python
"""profiles/traffic.json from the ncu launch lists tools/r2_traffic.sh wrote: DRAM bytes (read + write, summed over all
launches of one device-resident pass) per case, tagged with the kernel revision / mode / team of the plain run that preceded
the capture, so that bench.py drops an entry as stale when the kernels changed.
  python tools/update_traffic.py <tag> [dir with <tag>_launches_<w>.csv and <tag>_plain_<w>.log, default gpurun_out]"""
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def dram_bytes(path):
    tot, n = 0.0, set()
    for r in csv.reader(open(path)):
        if len(r) > 14 and r[0].isdigit() and r[12] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            tot += float(r[14].replace(",", "")) * UNIT.get(r[13], 1.0)
            n.add(int(r[0]))
    return tot, len(n)


def main(tag, d="gpurun_out"):
    rev = int(re.search(r"kKernelRev\s*=\s*(\d+)", open(os.path.join(ROOT, "fastbn_b200", "csrc", "capi.cu")).read()).group(1))
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    tj = json.load(open(tpath)) if os.path.exists(tpath) else {}
    for w in ["alarm", "hepar2", "hepar2_50", "pigs", "munin"]:
        lc, pl = os.path.join(ROOT, d, "%s_launches_%s.csv" % (tag, w)), os.path.join(ROOT, d, "%s_plain_%s.log" % (tag, w))
        if not (os.path.exists(lc) and os.path.exists(pl)):
            continue
        m = re.search(r"cases (\d+) mode (\d+) team (\d+)", open(pl).read())
        if not m:
            continue
        cases, mode, team = int(m.group(1)), int(m.group(2)), int(m.group(3))
        b, nl = dram_bytes(lc)
        if nl == 0:
            continue
        tj[w] = {"dram_bytes_per_case": b / cases, "kernel_rev": rev, "mode": mode, "team_warps": 0 if mode == 2 else team,
                 "source": "profiles/%s_launches_%s_summary.txt (ncu dram__bytes_read+write summed over the %d launches of one %d-case "
                           "device-resident pass, tools/r2_traffic.sh)" % (tag, w, nl, cases)}
        print(w, tj[w])
    json.dump(tj, open(tpath, "w"), indent=1)


def smem(workload, rep, cases, team):
    """--smem <workload> <file.ncu-rep> <cases> <team warps>: shared-memory bytes per case of the (single) profiled launch = LSU wavefronts x 128 B"""
    import subprocess
    rows = list(csv.reader(subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout.splitlines()))
    i = rows[0].index("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum")
    waves = float(rows[2][i].replace(",", ""))
    rev = int(re.search(r"kKernelRev\s*=\s*(\d+)", open(os.path.join(ROOT, "fastbn_b200", "csrc", "capi.cu")).read()).group(1))
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    tj = json.load(open(tpath))
    tj.setdefault(workload, {}).update({
        "smem_bytes_per_case": waves * 128.0 / cases, "smem_kernel_rev": rev, "smem_team_warps": team,
        "smem_source": "%s (ncu --set full, l1tex__data_pipe_lsu_wavefronts_mem_shared.sum x 128 B over one %d-case launch)" % (os.path.basename(rep), cases)})
    print(workload, tj[workload])
    json.dump(tj, open(tpath, "w"), indent=1)


if __name__ == "__main__":
    if sys.argv[1] == "--smem":
        smem(sys.argv[2], sys.argv[3], int(sys.argv[4]), int(sys.argv[5]))
    else:
        main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "gpurun_out")
