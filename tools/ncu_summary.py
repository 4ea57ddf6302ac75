#!/usr/bin/env python
"""Summarise an .ncu-rep (read in the no-GPU container): key raw metrics + per-source-line instruction/stall shares."""
import csv
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct',
        'l1tex__t_sector_hit_rate.pct', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__grid_size', 'launch__block_size', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_blocks',
        'smsp__cycles_active.avg', 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
        'smsp__warps_active.avg.per_cycle_active', 'smsp__warps_eligible.avg.per_cycle_active',
        # warp states per issued instruction (sum = resident warps per scheduler / issue rate)
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio']


def main(rep, top=25):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[0]
    print("== raw metrics (%s)" % rep)
    for k in KEYS:
        if k in hdr:
            i = hdr.index(k)
            print("%-70s %-12s %s" % (k, rows[1][i], " ".join(r[i] for r in rows[2:])))
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Line No')
    hdr = rows[hi]
    iS = hdr.index('# Samples')
    iI = hdr.index('Instructions Executed')
    lines = {}
    for r in rows[hi + 1:]:
        if r and r[0].strip().isdigit():
            try:
                ins, smp = int(r[iI]), int(r[iS])
            except ValueError:
                continue
            k = (int(r[0]), r[1][:100])
            a = lines.get(k, (0, 0))
            lines[k] = (a[0] + ins, a[1] + smp)
    tot = sum(v[0] for v in lines.values()) or 1
    ts = sum(v[1] for v in lines.values()) or 1
    print("== per source line: share of executed warp instructions / of stall samples")
    for (ln, src), (ins, smp) in sorted(lines.items(), key=lambda x: -x[1][0])[:top]:
        print("%4d inst %5.1f%% samp %5.1f%%  %s" % (ln, 100 * ins / tot, 100 * smp / ts, src))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25)
