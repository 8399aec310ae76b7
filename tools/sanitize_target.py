#!/usr/bin/env python
"""The program compute-sanitizer runs (memcheck, racecheck, initcheck): a 96x64 render of a 4,096-triangle mesh scene through every
kernel form -- device-built and host-built tree, wavefront with the loop on the device and on the host, the split form, wf_tail at a
low and at the default threshold, the megakernel, trace_rays, and a 2-member group on one device with both combines -- each
compared with the first result so that a run under the sanitizer is also a correctness run.

    compute-sanitizer --tool memcheck  python tools/sanitize_target.py
    compute-sanitizer --tool racecheck python tools/sanitize_target.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

from computeraytracer_b200 import scenes  # noqa: E402
from computeraytracer_b200.renderer import (BVH_BUILDER_DEVICE, BVH_BUILDER_HOST_SAH, KERNEL_MEGA, KERNEL_WAVEFRONT, SHARD_SPLIT_SPP, SHARD_TILES,
                                            WF_HOST_LOOP, WF_TAIL_OFF, Renderer, default_params)  # noqa: E402


def main():
    ps = scenes.mesh_scene(96, 64, n_lon=64, n_lat=32)
    ref = None
    with Renderer(0) as r:
        r.set_scene(ps)
        forms = [("wavefront, device loop, device tree", dict(kernel=KERNEL_WAVEFRONT, bvh_tuning=BVH_BUILDER_DEVICE, wf_slots=2048)),
                 ("wavefront, device loop, host tree, tail at 16", dict(kernel=KERNEL_WAVEFRONT, bvh_tuning=BVH_BUILDER_HOST_SAH, wf_slots=2048, wf_tuning=5 << 8)),
                 ("wavefront, device loop, two lanes", dict(kernel=KERNEL_WAVEFRONT, wf_slots=2048, wf_tuning=64 | (5 << 8))),
                 ("wavefront, host loop, no tail", dict(kernel=KERNEL_WAVEFRONT, wf_slots=4096, wf_tuning=WF_HOST_LOOP | WF_TAIL_OFF)),
                 ("wavefront, routed lists", dict(kernel=KERNEL_WAVEFRONT, wf_slots=4096, wf_tuning=2 | WF_TAIL_OFF)),
                 ("wavefront, split kernels", dict(kernel=KERNEL_WAVEFRONT, wf_slots=4096, flags=4)),
                 ("megakernel", dict(kernel=KERNEL_MEGA))]
        for name, kw in forms:
            r.set_params(default_params(max_depth=8, **kw)); r.reset()
            r.frame(1); r.frame(2)
            acc = r.accumulator()
            if ref is None:
                ref = acc
            assert np.array_equal(acc[..., :3].view(np.uint32), ref[..., :3].view(np.uint32)), name
            print("ok:", name, flush=True)
        lights = np.array([[[60, 500, 120], [180, 500, 120], [60, 500, 260]]], dtype=np.float32)
        r.set_scene(scenes.with_triangle_lights(ps, lights))
        tri_ref = None
        for name, kw in (("triangle light, wavefront", dict(kernel=KERNEL_WAVEFRONT, wf_slots=2048, wf_tuning=5 << 8)), ("triangle light, megakernel", dict(kernel=KERNEL_MEGA))):
            r.set_params(default_params(max_depth=8, **kw)); r.reset()
            r.frame(3)
            acc = r.accumulator()
            if tri_ref is None:
                tri_ref = acc
            assert np.array_equal(acc[..., :3].view(np.uint32), tri_ref[..., :3].view(np.uint32)), name
            print("ok:", name, flush=True)
        r.set_scene(ps); r.set_params(default_params(max_depth=8))
        rays = scenes.camera_rays(ps, 2048, seed=1)
        idx, t = r.trace_rays(rays)
        print("ok: trace_rays", int((idx >= 6).sum()), "triangle hits", flush=True)
    for mode in (SHARD_TILES, SHARD_SPLIT_SPP):
        with Renderer(devices=[0, 0], shard_mode=mode) as g:
            g.set_scene(ps); g.set_params(max_depth=8); g.reset()
            g.frame(3)
            acc = g.accumulator()
            d = np.abs(acc[..., :3].astype(np.float64) - ref[..., :3].astype(np.float64)) / 3
            assert d.max() <= 1e-3, mode
            if mode == SHARD_TILES:
                assert np.array_equal(acc[..., :3].view(np.uint32), ref[..., :3].view(np.uint32))
            print("ok: group of 2,", "tiles" if mode == SHARD_TILES else "split-spp", flush=True)
    print("sanitize target ok")


if __name__ == "__main__":
    main()
