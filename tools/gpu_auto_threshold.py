#!/usr/bin/env python
"""Tuning run (GPU box): time crt_dispatch_frames(n) of the reference's default scene with the megakernel and with the
wavefront form, for small n -- where AUTO should switch for linear-list scenes."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from computeraytracer_b200 import scenes                                   # noqa: E402
from computeraytracer_b200.renderer import Renderer, default_params        # noqa: E402

for (w, h) in ((1000, 1000), (1920, 1080)):
    ps = scenes.default_scene(w, h)
    with Renderer(0) as r:
        r.set_scene(ps)
        for n in (1, 2, 4, 8, 16, 64):
            row = []
            for kernel in (1, 2):
                r.set_params(default_params(kernel=kernel)); r.reset()
                r.frame(n); r.synchronize()
                reps = max(2, 32 // n)
                t0 = time.perf_counter()
                for _ in range(reps):
                    r.frame(n)
                r.synchronize()
                row.append((time.perf_counter() - t0) / reps * 1e3)
            print(f"{w}x{h} frames/dispatch {n:3d}: megakernel {row[0]:8.3f} ms  wavefront {row[1]:8.3f} ms  ratio {row[0] / row[1]:.2f}")
