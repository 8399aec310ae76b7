#!/usr/bin/env python
"""Diagnostics (GPU box): render a few samples of a scale scene with device statistics on and print the rays that
visit the most BVH nodes (CRT_SLOW_RAY_NODES threshold, logged by wf_traverse in stats builds).

    CRT_SLOW_RAY_NODES=1000 python tools/slow_rays.py [mesh1m|soup10m] [spp]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("CRT_SLOW_RAY_NODES", "1000")

from computeraytracer_b200 import scenes                      # noqa: E402
from computeraytracer_b200.renderer import Renderer, default_params   # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "mesh1m"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
ps = scenes.mesh_scene(1920, 1080) if which == "mesh1m" else scenes.soup_scene(1920, 1080)
with Renderer(0) as r:
    r.set_scene(ps)
    r.set_params(default_params(kernel=2, max_depth=100 if which == "mesh1m" else 8))
    r.enable_stats(True)
    r.reset(); r.reset_counters()
    r.frame(spp)
    c = r.counters()
    print({k: c[k] for k in ("paths", "rays_ext", "bvh_rays", "node_visits", "tri_tests", "max_ray_nodes", "wf_iterations")})
    print("nodes per BVH ray %.2f" % (c["node_visits"] / max(c["bvh_rays"], 1)))
