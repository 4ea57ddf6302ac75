#!/usr/bin/env python
"""Summarise an ncu launch list (csv of `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
--csv --log-file X`): per kernel name the launches, total time, DRAM bytes and GB/s; then the top launches.
  python tools/launch_summary.py gpurun_out/launches.csv [cases_in_the_run] [top]"""
import csv
import sys
from collections import defaultdict


def main(path, cases=0, top=6):
    rows = [r for r in csv.reader(open(path)) if len(r) > 14 and r[0].isdigit()]
    L = defaultdict(dict)
    for r in rows:
        L[int(r[0])]["name"] = r[4].split("(")[0][-40:]
        L[int(r[0])][r[12]] = float(r[14].replace(",", ""))
        L[int(r[0])]["unit_" + r[12]] = r[13]

    def ms(d):
        v, u = d.get("gpu__time_duration.sum", 0.0), d.get("unit_gpu__time_duration.sum", "ns")
        return v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1e-6)

    def gb(d):
        t = 0.0
        for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            v, u = d.get(k, 0.0), d.get("unit_" + k, "byte")
            t += v * {"byte": 1e-9, "Kbyte": 1e-6, "Mbyte": 1e-3, "Gbyte": 1.0}.get(u, 1e-9)
        return t

    tot_ms = sum(ms(d) for d in L.values())
    tot_gb = sum(gb(d) for d in L.values())
    line = "launches %d total ms %.2f dram GB %.2f avg GB/s %.0f" % (len(L), tot_ms, tot_gb, tot_gb / tot_ms * 1e3 if tot_ms else 0)
    if cases:
        line += " per case MB %.2f" % (tot_gb * 1e3 / cases)
    print(line)
    by = defaultdict(lambda: [0, 0.0, 0.0])
    for d in L.values():
        b = by[d["name"]]
        b[0] += 1; b[1] += ms(d); b[2] += gb(d)
    for n, b in sorted(by.items(), key=lambda x: -x[1][1]):
        print("%-42s n=%3d %8.2f ms %7.1f%% %8.2f GB %6.0f GB/s" % (n, b[0], b[1], 100 * b[1] / tot_ms, b[2], b[2] / b[1] * 1e3 if b[1] else 0))
    for i, d in sorted(L.items(), key=lambda x: -ms(x[1]))[:top]:
        print("top %3d %-40s %8.2f ms %8.2f GB %6.0f GB/s" % (i, d["name"], ms(d), gb(d), gb(d) / ms(d) * 1e3))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 0, int(sys.argv[3]) if len(sys.argv) > 3 else 6)
