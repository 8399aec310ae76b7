#!/usr/bin/env python
"""DRAM bytes per ray per kernel from an ncu launch list of tools/ncu_target.py (dram__bytes_read.sum + dram__bytes_write.sum
over exactly the profiled step's launches, divided by that step's rays) -> profiles/rNN_traffic.json, which bench.py reads
for roofline.traffic / dram_frac.   python tools/traffic_json.py <out.json> <workload>:<launches.csv>:<rays.json> ..."""
import collections
import csv
import json
import sys


def per_kernel(path):
    rows = list(csv.reader(open(path)))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    h = rows[hdr]
    ki, vi, mi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Name"), h.index("Metric Unit")
    agg = collections.defaultdict(lambda: collections.defaultdict(float))
    for r in rows[hdr + 1:]:
        if len(r) <= vi:
            continue
        name = r[ki].split("(")[0].split("<")[0].strip().split(" ")[-1]
        v = float(r[vi].replace(",", ""))
        if r[mi].startswith("dram__bytes"):
            agg[name]["dram"] += v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(r[ui], 1)
        elif r[mi] == "gpu__time_duration.sum":
            agg[name]["us"] += v * {"ns": 1e-3, "us": 1, "usecond": 1, "ms": 1e3, "msecond": 1e3}.get(r[ui], 1)
            agg[name]["n"] += 1
    return agg


def main():
    out = {"_comment": "DRAM traffic per ray of the hot-path kernels: ncu dram__bytes_read.sum + dram__bytes_write.sum over the launches of ONE "
                       "profiled step (tools/ncu_target.py, cudaProfilerStart/Stop around it) divided by that step's extension + shadow rays. "
                       "Under ncu every launch starts with a cold L2, so this is an upper bound for a warm run. bench.py scales it by the rays of its own run."}
    for spec in sys.argv[2:]:
        workload, csv_path, rays_path = spec.split(":")
        rays = json.load(open(rays_path))
        n_rays = rays["rays_ext"] + rays["rays_shadow"]
        out[workload] = {"_step": rays}
        for k, a in per_kernel(csv_path).items():
            if a["n"]:
                out[workload][k] = {"dram_bytes_per_ray": a["dram"] / max(n_rays, 1), "launches": int(a["n"]), "time_us": a["us"]}
    json.dump(out, open(sys.argv[1], "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
