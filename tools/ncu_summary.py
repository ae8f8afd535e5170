#!/usr/bin/env python
"""Print the handful of `ncu --page raw` counters DESIGN.md argues from, one block per profiled launch.

usage: tools/ncu_summary.py <report.ncu-rep>
"""
import csv
import io
import subprocess
import sys

WANT = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "sass__inst_executed_local_loads", "sass__inst_executed_local_stores"]


def main():
    if sys.argv[1].endswith(".csv"):      # `ncu -i <rep> --page raw --csv`, exported on the GPU box (tools/ncu_export.sh)
        raw = open(sys.argv[1]).read()
    else:
        raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units = rows[0], rows[1]
    stalls = [h for h in head if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
    for row in rows[2:]:
        print("=" * 100)
        for w in WANT:
            if w in head:
                i = head.index(w)
                print(f"{w} [{units[i]}] = {row[i]}")
        st = sorted(((float(row[head.index(s)] or 0), s) for s in stalls), reverse=True)[:7]
        for v, s in st:
            print(f"  stall {s[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]:24s} {v:.3f} warps per issue")


if __name__ == "__main__":
    main()
