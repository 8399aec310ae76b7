#!/usr/bin/env python
"""Diagnostic for tests/test_wgsl_literal.py's random scenes: libcrt (a kernel form) against the oracle, per seed the number of
pixels that differ and a few of them.    python tools/fuzz_diag.py [n_seeds [first_seed]]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from computeraytracer_b200.renderer import KERNEL_MEGA, KERNEL_WAVEFRONT, WF_HOST_LOOP, WF_TAIL_OFF, Renderer  # noqa: E402
from tests import orc  # noqa: E402
from tests.test_wgsl_literal import _random_scene_with_triangles  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    first = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    o = orc.load()
    W = H = 40
    forms = [("mega", KERNEL_MEGA, 0), ("wavefront", KERNEL_WAVEFRONT, 0), ("wf_host_loop_no_tail", KERNEL_WAVEFRONT, WF_HOST_LOOP | WF_TAIL_OFF)]
    for seed in range(first, first + n):
        ps = _random_scene_with_triangles(seed, W, H)
        acc_c, fb_c, cc = o.scene(ps, orc.default_params(max_depth=12)).render(3)
        for name, k, tune in forms:
            with Renderer(0) as r:
                r.set_params(kernel=k, max_depth=12, wf_tuning=tune)
                r.set_scene(ps)
                r.frame(1); r.frame(2)
                acc, fb, c = r.accumulator(), r.framebuffer(), r.counters()
            a, b = acc[..., :3], acc_c[..., :3]
            same = ((a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))).all(axis=-1)
            bad = np.argwhere(~same)
            print(f"seed {seed} {name}: {len(bad)} pixels differ, fb equal {np.array_equal(fb, fb_c)}, rays {c['rays_ext']}/{int(cc[1])} {c['rays_shadow']}/{int(cc[2])}")
            for y, x in bad[:4]:
                print("    ", (int(x), int(y)), "gpu", a[y, x], "oracle", b[y, x])


if __name__ == "__main__":
    main()
