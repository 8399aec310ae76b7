"""The oracle's own BVH path (used for full-frame parity on the big scenes) returns exactly what
its linear scan returns -- hit index and t, bit for bit -- including exact-t ties and excludes."""
import numpy as np
import pytest

from computeraytracer_b200 import scenes


def _cmp(oracle, ps, rays, exclude=None):
    lin = oracle.scene(ps)
    acc = oracle.scene(ps, use_bvh=True)
    i0, t0, _ = lin.trace_rays(rays, exclude)
    i1, t1, c1 = acc.trace_rays(rays, exclude)
    assert np.array_equal(i0, i1)
    assert np.array_equal(t0.view(np.uint32), t1.view(np.uint32))
    return i0, c1


def test_mesh_scene_bvh_equals_linear(oracle):
    ps = scenes.mesh_scene(64, 64, n_lon=64, n_lat=32)
    rays = scenes.camera_rays(ps, 4000, seed=1)
    idx, cnt = _cmp(oracle, ps, rays)
    assert (idx >= 6).sum() > 200 and (idx < 6).sum() > 200
    assert cnt[5] < 4000 * 4096 / 4  # the BVH actually culls


def test_soup_scene_bvh_equals_linear(oracle):
    ps = scenes.soup_scene(64, 64, n=3000)
    rays = scenes.camera_rays(ps, 4000, seed=2)
    rng = np.random.default_rng(0)
    exclude = rng.integers(0, 3006, size=4000).astype(np.uint32)
    _cmp(oracle, ps, rays)
    _cmp(oracle, ps, rays, exclude)


def test_exact_ties_go_to_the_highest_index(oracle):
    """Two coincident triangles and a triangle coplanar with a patch: the later primitive wins (CS:557)."""
    ps = scenes.default_scene(32, 32)
    tri = np.array([[[100, 100, 300], [200, 100, 300], [100, 200, 300]]] * 3, dtype=np.float32)
    # third copy lies in the back wall's plane z = 555 (patch 3)
    tri[2, :, 2] = 555.0
    ps2 = scenes._with_triangles(ps, 6, tri)
    rays = np.zeros((2, 8), dtype=np.float32)
    rays[:, 0:3] = (120, 120, 0)
    rays[:, 4:7] = (0, 0, 1)
    idx, _ = _cmp(oracle, ps2, rays)
    assert idx[0] == 7  # of the coincident pair 6, 7 the later one
    ex = np.array([7, 6], dtype=np.uint32)
    idx, _ = _cmp(oracle, ps2, rays, ex)
    assert idx.tolist() == [6, 7]
    ex = np.array([6, 6], dtype=np.uint32)
    rays[:, 0:3] = (120, 120, 400)  # between the pair and the wall: wall (3) vs coplanar triangle (8)
    idx, _ = _cmp(oracle, ps2, rays, ex)
    assert idx.tolist() == [8, 8]


def test_full_frame_bvh_equals_linear(oracle):
    ps = scenes.mesh_scene(48, 48, n_lon=32, n_lat=16)
    a0, f0, c0 = oracle.scene(ps).render(2)
    a1, f1, c1 = oracle.scene(ps, use_bvh=True).render(2)
    assert np.array_equal(a0.view(np.uint32), a1.view(np.uint32))
    assert np.array_equal(f0, f1)
    assert c0[0] == c1[0] and c0[1] == c1[1] and c0[2] == c1[2]
