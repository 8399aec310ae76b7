#!/usr/bin/env python
"""The program ncu profiles: one step of a workload between cudaProfilerStart / Stop (run ncu with --profile-from-start off),
host-driven iteration loop so that every kernel is its own launch.  Writes the step's ray counts next to the ncu output so
that DRAM bytes per ray can be computed for exactly the profiled launches (tools/traffic_json.py).

    python tools/ncu_target.py <workload> <spp> <rays.json> [--graph]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from bench import workload_scene  # noqa: E402
from computeraytracer_b200.renderer import WF_HOST_LOOP, Renderer  # noqa: E402


def main():
    workload, spp, out = sys.argv[1], int(sys.argv[2]), sys.argv[3]
    ps, max_depth, _ = workload_scene(workload, 1920, 1080)
    r = Renderer(0)
    r.set_params(max_depth=max_depth, wf_tuning=0 if "--graph" in sys.argv else WF_HOST_LOOP)
    r.set_scene(ps)
    r.frame(spp); r.synchronize()          # warm-up: allocations, scene upload
    r.reset(); r.reset_counters()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
    flush.zero_(); torch.cuda.synchronize()
    torch.cuda.profiler.start()
    r.frame(spp); r.synchronize()
    torch.cuda.profiler.stop()
    c = r.counters()
    json.dump({"workload": workload, "spp": spp, "rays_ext": c["rays_ext"], "rays_shadow": c["rays_shadow"], "paths": c["paths"],
               "launches": c["launches"]}, open(out, "w"))
    r.close()


if __name__ == "__main__":
    main()
