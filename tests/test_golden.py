"""Committed golden vectors (tests/golden/cornell_oracle_golden.npz, made by tests/golden/make_fixtures.py).

The reference has no tests or fixtures for this path and its WGSL cannot run here, so the goldens are the
oracle's own outputs on the reference's default scene, frozen after the oracle was checked against the
survey-derived known answers (test_oracle_kat.py).  CPU: the oracle still reproduces them bit for bit.
GPU: the CUDA path (through the C ABI) reproduces them without the oracle in the loop.
"""
import os

import numpy as np
import pytest

from computeraytracer_b200 import scenes

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cornell_oracle_golden.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def bits(a):
    return np.ascontiguousarray(a[..., :3]).view(np.uint32)


def test_oracle_reproduces_golden_frames(oracle, gold):
    from tests import orc
    acc, fb, cnt = oracle.scene(scenes.default_scene(64, 64)).render(4)
    assert np.array_equal(bits(acc), bits(gold["acc_64x64_s4"]))
    assert np.array_equal(fb, gold["fb_64x64_s4"])
    assert np.array_equal(cnt, gold["cnt_64x64_s4"])
    acc, fb, cnt = oracle.scene(scenes.default_scene(50, 30), orc.default_params(max_depth=4)).render(3)
    assert np.array_equal(bits(acc), bits(gold["acc_50x30_s3_d4"]))
    assert np.array_equal(fb, gold["fb_50x30_s3_d4"])
    assert np.array_equal(cnt, gold["cnt_50x30_s3_d4"])


def test_oracle_reproduces_golden_hits(oracle, gold):
    sc = oracle.scene(scenes.default_scene(64, 64))
    i0, t0, _ = sc.trace_rays(gold["rays"], None)
    i1, t1, _ = sc.trace_rays(gold["rays"], gold["exclude"])
    assert np.array_equal(i0, gold["hit_idx"]) and np.array_equal(t0.view(np.uint32), gold["hit_t"].view(np.uint32))
    assert np.array_equal(i1, gold["hit_idx_ex"]) and np.array_equal(t1.view(np.uint32), gold["hit_t_ex"].view(np.uint32))
    assert len(np.unique(i0)) >= 15   # the ray set sees nearly every primitive


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", [1, 2])   # CRT_KERNEL_MEGA, CRT_KERNEL_WAVEFRONT
def test_gpu_reproduces_golden_frames(gold, kernel):
    from computeraytracer_b200.renderer import Renderer
    with Renderer(0) as r:
        r.set_scene(scenes.default_scene(64, 64))
        r.set_params(kernel=kernel)
        r.frame(4)
        assert np.array_equal(bits(r.accumulator()), bits(gold["acc_64x64_s4"]))
        assert np.array_equal(r.framebuffer(), gold["fb_64x64_s4"])
        c = r.counters()
        assert [c["paths"], c["rays_ext"], c["rays_shadow"]] == [int(x) for x in gold["cnt_64x64_s4"][:3]]
        r.set_scene(scenes.default_scene(50, 30))
        r.set_params(max_depth=4, kernel=kernel)
        r.reset()
        r.frame(3)
        assert np.array_equal(bits(r.accumulator()), bits(gold["acc_50x30_s3_d4"]))
        assert np.array_equal(r.framebuffer(), gold["fb_50x30_s3_d4"])


@pytest.mark.gpu
def test_gpu_reproduces_golden_hits(gold):
    from computeraytracer_b200.renderer import Renderer
    with Renderer(0) as r:
        r.set_scene(scenes.default_scene(64, 64))
        i0, t0 = r.trace_rays(gold["rays"], None)
        i1, t1 = r.trace_rays(gold["rays"], gold["exclude"])
    assert np.array_equal(i0, gold["hit_idx"]) and np.array_equal(t0.view(np.uint32), gold["hit_t"].view(np.uint32))
    assert np.array_equal(i1, gold["hit_idx_ex"]) and np.array_equal(t1.view(np.uint32), gold["hit_t_ex"].view(np.uint32))
