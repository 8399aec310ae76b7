#!/usr/bin/env python
"""Summarise an `ncu --metrics ... --csv` launch list: per kernel count, total time, time-weighted metrics.
    python tools/launch_summary.py launches.csv [--seq N]"""
import collections
import csv
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    h = rows[hdr]
    ki, vi, mi, ii, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Name"), h.index("ID"), h.index("Metric Unit")
    per = {}
    for r in rows[hdr + 1:]:
        if len(r) <= vi:
            continue
        v = float(r[vi].replace(",", ""))
        if r[mi].startswith("dram__bytes"):
            v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(r[ui], 1)
        if r[mi] == "gpu__time_duration.sum":
            v *= {"ns": 1e-3, "us": 1, "usecond": 1, "ms": 1e3, "msecond": 1e3}.get(r[ui], 1)
        per.setdefault(int(r[ii]), {"k": r[ki].split("(")[0][-28:]})[r[mi]] = v
    agg = collections.defaultdict(lambda: collections.defaultdict(float))
    for i in sorted(per):
        p = per[i]
        t = p.get("gpu__time_duration.sum", 0.0)
        a = agg[p["k"]]
        a["n"] += 1
        a["t"] += t
        a["dram"] += p.get("dram__bytes_read.sum", 0) + p.get("dram__bytes_write.sum", 0)
        for m, short in (("smsp__thread_inst_executed_per_inst_executed.ratio", "thr"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ"),
                         ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue")):
            a[short] += p.get(m, 0) * t
    tot = sum(a["t"] for a in agg.values())
    print(f"{'kernel':30s} {'n':>6s} {'time_us':>12s} {'share':>6s} {'thr/inst':>8s} {'occ%':>6s} {'issue%':>6s} {'dram_GB':>9s} {'GB/s':>8s}")
    for k, a in sorted(agg.items(), key=lambda x: -x[1]["t"]):
        t = max(a["t"], 1e-9)
        print(f"{k:30s} {int(a['n']):6d} {a['t']:12.1f} {100 * a['t'] / tot:5.1f}% {a['thr'] / t:8.1f} {a['occ'] / t:6.1f} {a['issue'] / t:6.1f} {a['dram'] / 1e9:9.2f} {a['dram'] / t / 1e3:8.1f}")
    if "--seq" in sys.argv:
        n = int(sys.argv[sys.argv.index("--seq") + 1])
        for i in sorted(per)[:n]:
            p = per[i]
            print(i, p["k"], round(p.get("gpu__time_duration.sum", 0), 1), round(p.get("smsp__thread_inst_executed_per_inst_executed.ratio", 0), 1))


if __name__ == "__main__":
    main()
