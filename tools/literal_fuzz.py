#!/usr/bin/env python
"""CPU only: the C oracle against the Python reading of the shader (tests/wgsl_literal.py) on random scenes
(tests/test_wgsl_literal.py::_random_scene), 16x16 x 2 frames each: accumulator bits (NaN = NaN), statistics, framebuffer bytes.
    python tools/literal_fuzz.py FIRST_SEED LAST_SEED      (the CPU suite runs seeds 2, 3, 4, 7; round 2 ran 0..99: all agree)"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import orc, wgsl_literal as wl  # noqa: E402
from tests.test_wgsl_literal import _random_scene, _same_bits_or_both_nan, _statistics  # noqa: E402


def main():
    o = orc.load()
    part = lambda i: (lambda x: o.detmath(np.array([x], dtype=np.float32))[i][0])  # noqa: E731
    sincos = lambda x: (part(0)(x), part(1)(x))  # noqa: E731
    W = H = 16
    bad = 0
    for seed in range(int(sys.argv[1]), int(sys.argv[2])):
        ps = _random_scene(seed, W, H)
        sh = wl.Shader(ps.bindings(), wl.Machine(sincos, part(2), part(3), fused=True))
        acc_l = np.zeros((H, W, 3), np.float32)
        for s in (1, 2):
            for y in range(H):
                for x in range(W):
                    acc_l[y, x] += np.array(sh.main(x, y, s), dtype=np.float32)
        acc, fb, cnt = o.scene(ps).render(2, threads=1)
        fb_l = np.array([[sh.resolve(acc_l[y, x], 2) for x in range(W)] for y in range(H)], dtype=np.uint8)
        ok = (_same_bits_or_both_nan(acc_l, acc[..., :3]), _statistics(sh) == [int(c) for c in cnt[:12]], bool(np.array_equal(fb_l, fb)))
        bad += not all(ok)
        print(seed, "prims", ps.primitives.shape[0], "lights", ps.lights.shape[0], "acc/stats/fb", ok, "nan", int(np.isnan(acc_l).sum()), sh.term, sh.events)
    print("disagreements:", bad)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
