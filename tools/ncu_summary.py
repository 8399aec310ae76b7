#!/usr/bin/env python
"""Key metrics of an `ncu --set full` report, one block per captured launch.
    python tools/ncu_summary.py REPORT.ncu-rep [more reports...]"""
import csv
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_active",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
]
STALL = "smsp__average_warps_issue_stalled_"


def main():
    for rep in sys.argv[1:]:
        txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(txt.splitlines()))
        h, units = rows[0], rows[1]
        for r in rows[2:]:
            print(f"=== {rep}  launch {r[h.index('ID')]}  {r[h.index('Kernel Name')][:60]}")
            for k in KEYS:
                if k in h:
                    print(f"  {k:72s} {r[h.index(k)]:>16s} {units[h.index(k)]}")
            st = [(float(r[i] or 0), h[i][len(STALL):].replace("_per_issue_active.ratio", "")) for i in range(len(h))
                  if h[i].startswith(STALL) and h[i].endswith("_per_issue_active.ratio")]
            print("  stalls (warps per issue): " + ", ".join(f"{n} {v:.2f}" for v, n in sorted(st, reverse=True)[:7]))


if __name__ == "__main__":
    main()
