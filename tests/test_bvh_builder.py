"""The scene-upload module's BVH builders checked on the CPU through the host-only hook crt_host_build_bvh: structure of
the Wide8 tree, conservative quantised boxes, determinism, edge cases.  Two builders: the host binned-SAH build
(csrc/crt_bvh.cpp) and the device build's algorithm (csrc/crt_bvh_ploc.h: Morton sort, PLOC, collapse, level-by-level
emission) -- the CUDA kernels of csrc/crt_bvh_device.cu call the same per-element functions with one thread per index, and
here a host loop drives them (builder = DEVICE), so that logic is proven without a GPU; the -m gpu tests then check the tree
the kernels actually leave on the device (crt_read_bvh) with the same walk.

The tree only decides which triangles get tested (DESIGN.md "Conservative culling"): a triangle that is missing from the
leaves, or a child box that does not contain its triangles, would make the CUDA traversal drop hits the reference's
linear scan (ComputeShader.wgsl:503-518) finds.  The GPU parity tests catch that on the scenes they render; this test
proves the invariants directly, node by node, for every tree it builds.
"""
import ctypes as C

import numpy as np
import pytest

from computeraytracer_b200 import scenes
from computeraytracer_b200.scene_loader import PRIM_DTYPE

WORDS = 20  # 5 x 16 bytes per node


@pytest.fixture(scope="module")
def lib():
    from computeraytracer_b200.build import build_libcrt
    lib = C.CDLL(build_libcrt())
    lib.crt_host_build_bvh.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                       C.c_void_p, C.c_void_p, C.c_uint32, C.c_float]
    lib.crt_host_build_bvh.restype = C.c_int
    return lib


HOST_SAH, DEVICE_PLOC, EXACT_SLOTS = 0x400, 0x800, 0x200   # crt_params.bvh_tuning bits (include/crt.h)
BUILDERS = [pytest.param(HOST_SAH, id="host_sah"), pytest.param(DEVICE_PLOC, id="ploc")]


def build(lib, prims: np.ndarray, threads: int = 4, tuning: int = 0, prim_cost: float = 0.0):
    raw = np.ascontiguousarray(prims).view(np.uint8)
    info = np.zeros(4, dtype=np.uint64)
    bounds = np.zeros(7, dtype=np.float32)
    assert lib.crt_host_build_bvh(raw.ctypes.data, raw.size, threads, None, 0, None, 0, info.ctypes.data, bounds.ctypes.data, tuning, prim_cost) == 0
    n_nodes, n_tri, depth, words = (int(x) for x in info)
    assert words == WORDS
    nodes = np.zeros(max(n_nodes, 1) * WORDS, dtype=np.uint32)
    src = np.zeros(max(n_tri, 1), dtype=np.uint32)
    assert lib.crt_host_build_bvh(raw.ctypes.data, raw.size, threads, nodes.ctypes.data, nodes.size, src.ctypes.data, src.size,
                                  info.ctypes.data, bounds.ctypes.data, tuning, prim_cost) == 0
    return nodes[: n_nodes * WORDS].reshape(n_nodes, WORDS), src[:n_tri], depth, bounds


def tri_prims(tris: np.ndarray, extra_front: int = 0) -> np.ndarray:
    """binding-8 records: `extra_front` patches (category 0, ignored by the builder), then the triangles."""
    n = tris.shape[0]
    p = np.zeros(extra_front + n, dtype=PRIM_DTYPE)
    sl = slice(extra_front, extra_front + n)
    t = np.asarray(tris, dtype="<f4")
    p["category"][sl] = 2
    p["data1"][sl] = t[:, 0]
    p["data2"][sl] = t[:, 1] - t[:, 0]
    p["data3"][sl] = t[:, 2] - t[:, 0]
    p["data4"][sl, 3] = np.arange(extra_front, extra_front + n, dtype=np.uint32)
    return p


def tri_boxes(prims: np.ndarray):
    """per record: min / max corner of (v0, v0 + e1, v0 + e2), evaluated in f32 like the builder does"""
    v0 = prims["data1"].astype(np.float32)
    v1 = (v0 + prims["data2"].astype(np.float32)).astype(np.float32)
    v2 = (v0 + prims["data3"].astype(np.float32)).astype(np.float32)
    lo = np.minimum(v0, np.minimum(v1, v2))
    hi = np.maximum(v0, np.maximum(v1, v2))
    return lo, hi


def check_tree(nodes: np.ndarray, src: np.ndarray, depth: int, bounds: np.ndarray, prims: np.ndarray):
    """Walks the whole tree.  Returns (triangles seen, leaves)."""
    n_tri = int((prims["category"] == 2).sum())
    assert src.size == n_tri
    assert sorted(src.tolist()) == np.nonzero(prims["category"] == 2)[0].tolist()   # every triangle exactly once
    if n_tri == 0:
        assert nodes.shape[0] == 0
        return 0, 0
    lo, hi = tri_boxes(prims)
    pad = float(bounds[6])
    assert pad > 0.0
    # the root box holds every triangle with the padding to spare
    assert np.all(bounds[0:3] <= lo[src].min(axis=0) - pad * 0.999) and np.all(bounds[3:6] >= hi[src].max(axis=0) + pad * 0.999)
    seen_slots = np.zeros(n_tri, dtype=np.int32)
    visited = np.zeros(nodes.shape[0], dtype=np.int32)
    leaves = 0
    max_depth = 0

    def subtree(k: int, level: int):
        """(lo, hi) of all triangles below node k, checked against every child box on the way up"""
        nonlocal leaves, max_depth
        visited[k] += 1
        max_depth = max(max_depth, level)
        w = nodes[k]
        p = w[0:3].view(np.float32).astype(np.float64)
        e = [(int(w[3]) >> (8 * a)) & 0xFF for a in range(3)]
        imask = int(w[3]) >> 24
        step = np.array([2.0 ** (x - 127 - 15) for x in e])       # the node stores the exponent of 2^15 * step
        child_base, tri_base = int(w[4]), int(w[5])
        meta = w[6:8].view(np.uint8)
        qlo = np.stack([w[8:10].view(np.uint8), w[10:12].view(np.uint8), w[12:14].view(np.uint8)])   # [axis][slot]
        qhi = np.stack([w[14:16].view(np.uint8), w[16:18].view(np.uint8), w[18:20].view(np.uint8)])
        tot_lo = np.full(3, np.inf); tot_hi = np.full(3, -np.inf)
        n_children = 0
        for sl in range(8):
            m = int(meta[sl])
            if m == 0:
                assert not (imask >> sl) & 1
                continue
            n_children += 1
            if (m >> 5) == 1 and (m & 31) >= 24:                  # inner child: 001 11sss
                assert (m & 31) == 24 + sl and (imask >> sl) & 1
                child = child_base + bin(imask & ((1 << sl) - 1)).count("1")
                assert k < child < nodes.shape[0]                 # breadth-first numbering: children after their parent
                c_lo, c_hi = subtree(child, level + 1)
            else:                                                 # leaf: unary count << 5 | offset
                assert not (imask >> sl) & 1
                count = {1: 1, 3: 2, 7: 3}[m >> 5]
                off = m & 31
                assert off + count <= 24
                slots = np.arange(tri_base + off, tri_base + off + count)
                assert slots[-1] < n_tri
                seen_slots[slots] += 1
                leaves += 1
                ids = src[slots]
                c_lo, c_hi = lo[ids].min(axis=0).astype(np.float64), hi[ids].max(axis=0).astype(np.float64)
            # conservative: the dequantised box holds the child's triangles with the padding
            b_lo = p + qlo[:, sl].astype(np.float64) * step
            b_hi = p + qhi[:, sl].astype(np.float64) * step
            # (the builder pads in f32: lo - pad and hi + pad are rounded once, hence the ulp of slack)
            ulp = np.spacing(np.maximum(np.abs(c_lo), np.abs(c_hi)).astype(np.float32)).astype(np.float64)
            assert np.all(b_lo <= c_lo - pad * 0.999 + ulp), (k, sl, b_lo, c_lo, pad)
            assert np.all(b_hi >= c_hi + pad * 0.999 - ulp), (k, sl, b_hi, c_hi, pad)
            tot_lo = np.minimum(tot_lo, c_lo); tot_hi = np.maximum(tot_hi, c_hi)
        assert n_children >= 1
        return tot_lo, tot_hi

    import sys
    sys.setrecursionlimit(10000)
    subtree(0, 1)
    assert np.all(visited == 1)            # a tree: every node has exactly one parent, none is orphaned
    assert np.all(seen_slots == 1)         # every triangle slot belongs to exactly one leaf
    assert max_depth == depth
    return n_tri, leaves


@pytest.mark.parametrize("builder", BUILDERS)
def test_mesh_tree_is_complete_and_conservative(lib, builder):
    ps = scenes.mesh_scene(64, 48, n_lon=96, n_lat=48)           # 9,216 triangles behind 6 patches
    nodes, src, depth, bounds = build(lib, ps.primitives, tuning=builder)
    n_tri, leaves = check_tree(nodes, src, depth, bounds, ps.primitives)
    assert n_tri == 2 * 96 * 48 and leaves >= n_tri // 3
    assert 3 <= depth <= 12


@pytest.mark.parametrize("builder", BUILDERS)
def test_soup_tree_is_complete_and_conservative(lib, builder):
    ps = scenes.soup_scene(64, 48, n=20_000)
    nodes, src, depth, bounds = build(lib, ps.primitives, tuning=builder)
    assert check_tree(nodes, src, depth, bounds, ps.primitives)[0] == 20_000


def test_ploc_build_is_deterministic_and_as_good_as_the_sah_build(lib):
    """The device algorithm gives the same bytes every time (scans, no atomics), and its trees are in the SAH build's
    class: summed area of all child boxes (what a random ray pays for) within 25 % on a mesh and on a soup."""
    def child_area(nodes):
        tot = 0.0
        for w in nodes:
            e = [(int(w[3]) >> (8 * a)) & 0xFF for a in range(3)]
            step = np.array([2.0 ** (x - 127 - 15) for x in e])
            meta = w[6:8].view(np.uint8)
            qlo = np.stack([w[8:10].view(np.uint8), w[10:12].view(np.uint8), w[12:14].view(np.uint8)]).astype(np.float64)
            qhi = np.stack([w[14:16].view(np.uint8), w[16:18].view(np.uint8), w[18:20].view(np.uint8)]).astype(np.float64)
            d = (qhi - qlo) * step[:, None]
            a = d[0] * d[1] + d[1] * d[2] + d[2] * d[0]
            tot += float(a[meta != 0].sum())
        return tot
    for ps in (scenes.mesh_scene(64, 48, n_lon=96, n_lat=48), scenes.soup_scene(64, 48, n=20_000)):
        a = build(lib, ps.primitives, tuning=DEVICE_PLOC)
        b = build(lib, ps.primitives, tuning=DEVICE_PLOC)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[2] == b[2]
        sah = build(lib, ps.primitives, tuning=HOST_SAH)
        assert child_area(a[0]) <= 1.25 * child_area(sah[0]), (child_area(a[0]), child_area(sah[0]))


@pytest.mark.parametrize("threads", [1, 2, 7, 16])
def test_result_does_not_depend_on_the_thread_count(lib, threads):
    ps = scenes.soup_scene(64, 48, n=30_000)
    ref = build(lib, ps.primitives, 1)
    got = build(lib, ps.primitives, threads)
    assert np.array_equal(ref[0], got[0]) and np.array_equal(ref[1], got[1]) and ref[2] == got[2]
    assert np.array_equal(ref[3], got[3])


@pytest.mark.parametrize("builder", BUILDERS)
def test_no_triangles_and_single_triangle(lib, builder):
    ps = scenes.default_scene(32, 32)                            # patches and spheres only
    nodes, src, depth, bounds = build(lib, ps.primitives, tuning=builder)
    assert nodes.shape[0] == 0 and src.size == 0
    one = tri_prims(np.array([[[1, 2, 3], [4, 2, 3], [1, 6, 3]]], dtype=np.float32), extra_front=3)
    nodes, src, depth, bounds = build(lib, one, tuning=builder)
    assert nodes.shape[0] == 1 and src.tolist() == [3] and depth == 1
    check_tree(nodes, src, depth, bounds, one)
    empty = np.zeros(0, dtype=PRIM_DTYPE)
    nodes, src, depth, bounds = build(lib, empty, tuning=builder)
    assert nodes.shape[0] == 0 and src.size == 0


@pytest.mark.parametrize("builder", BUILDERS)
@pytest.mark.parametrize("n", [2, 3, 4, 9, 25, 200])
def test_small_counts(lib, n, builder):
    rng = np.random.default_rng(n)
    tris = rng.uniform(0, 100, size=(n, 3, 3)).astype(np.float32)
    p = tri_prims(tris)
    nodes, src, depth, bounds = build(lib, p, tuning=builder)
    assert check_tree(nodes, src, depth, bounds, p)[0] == n


@pytest.mark.parametrize("builder", BUILDERS)
def test_coincident_centroids_and_degenerate_triangles(lib, builder):
    """All centroids equal (no SAH split exists: the builder falls back to object-median splits), zero-area triangles,
    exact duplicates and a very flat scene (one axis has no extent)."""
    base = np.array([[0, 0, 0], [3, 0, 0], [0, 3, 0]], dtype=np.float32) - np.array([1, 1, 0], dtype=np.float32)
    same = np.repeat(base[None], 300, axis=0) * np.linspace(1, 2, 300, dtype=np.float32)[:, None, None]   # centroids all (0,0,0)
    dup = np.repeat(base[None] + 10.0, 64, axis=0)
    point = np.zeros((50, 3, 3), dtype=np.float32) + 7.0          # degenerate: all three vertices equal
    rng = np.random.default_rng(3)
    flat = rng.uniform(0, 50, size=(500, 3, 3)).astype(np.float32); flat[:, :, 2] = 5.0
    for tris in (same, dup, point, flat, np.concatenate([same, dup, point, flat])):
        p = tri_prims(tris, extra_front=2)
        nodes, src, depth, bounds = build(lib, p, tuning=builder)
        assert check_tree(nodes, src, depth, bounds, p)[0] == tris.shape[0]
        # CRT_BVH_STACK: the traversal stack holds 32 node groups.  The SAH build bounds its depth (median splits); the
        # agglomerative build turns 300 nested triangles into a chain (depth 43 here) -- a valid tree that upload_scene
        # does not use: it falls back to the SAH build when the device tree is deeper than the stack (tested on the GPU)
        if builder == HOST_SAH:
            assert depth + 2 <= 32


@pytest.mark.parametrize("builder", BUILDERS)
def test_large_coordinates_and_tiny_extents(lib, builder):
    """The quantisation frame (origin + power-of-two step per axis) stays conservative far from the origin and for
    nodes much smaller than the scene."""
    rng = np.random.default_rng(11)
    c = rng.uniform(-1, 1, size=(4000, 1, 3)) * np.array([1e5, 1e-2, 10.0])
    tris = (c + rng.uniform(-1e-3, 1e-3, size=(4000, 3, 3)) + np.array([3e5, -7e3, 0.0])).astype(np.float32)
    p = tri_prims(tris)
    nodes, src, depth, bounds = build(lib, p, tuning=builder)
    assert check_tree(nodes, src, depth, bounds, p)[0] == 4000


@pytest.mark.parametrize("builder", BUILDERS)
def test_padding_follows_the_coordinate_magnitude(lib, builder):
    """A mesh of extent ~2 translated to x = 4000: f32 rounds its vertices (v0 + e1) and every o - v0 to an ulp of the
    COORDINATE (2.4e-4 there), so a pad tied to the extent alone (2 / 32768 = 6e-5) would be below the rounding it has to
    cover.  The pad must scale with max |coordinate| as well."""
    rng = np.random.default_rng(5)
    tris = (rng.uniform(-1, 1, size=(3000, 1, 3)) + rng.uniform(-0.02, 0.02, size=(3000, 3, 3))).astype(np.float32)
    near = build(lib, tri_prims(tris), tuning=builder)
    far_tris = (tris.astype(np.float64) + np.array([4000.0, 0.0, 0.0])).astype(np.float32)
    p = tri_prims(far_tris)
    nodes, src, depth, bounds = build(lib, p, tuning=builder)
    assert check_tree(nodes, src, depth, bounds, p)[0] == 3000
    assert float(near[3][6]) == pytest.approx(max(2.04, 1.02) / 32768.0, rel=0.05)     # extent rules near the origin
    assert float(bounds[6]) >= 4000.0 / 32768.0                                         # the coordinate rules far from it
    assert float(bounds[6]) >= 256 * float(np.spacing(np.float32(4001.0)))


def test_argument_errors(lib):
    info = np.zeros(4, dtype=np.uint64)
    one = tri_prims(np.array([[[1, 2, 3], [4, 2, 3], [1, 6, 3]]], dtype=np.float32))
    raw = one.view(np.uint8)
    assert lib.crt_host_build_bvh(raw.ctypes.data, 79, 1, None, 0, None, 0, info.ctypes.data, None, 0, 0.0) == 1      # CRT_ERR_ARG: not 80-byte records
    assert lib.crt_host_build_bvh(raw.ctypes.data, 80, 1, None, 0, None, 0, None, None, 0, 0.0) == 1
    nodes = np.zeros(4, dtype=np.uint32); src = np.zeros(1, dtype=np.uint32)
    assert lib.crt_host_build_bvh(raw.ctypes.data, 80, 1, nodes.ctypes.data, nodes.size, src.ctypes.data, 1, info.ctypes.data, None, 0, 0.0) == 3   # CRT_ERR_SIZE


@pytest.mark.parametrize("bins", [8, 128, 256])
def test_bin_count_override_builds_a_valid_tree(lib, bins):
    """crt_params.bvh_tuning bits 0-8 (SAH bins; tools/bvh_eval.py) change the shape of the tree, never its invariants."""
    ps = scenes.soup_scene(64, 48, n=20_000)
    ref = build(lib, ps.primitives)
    nodes, src, depth, bounds = build(lib, ps.primitives, tuning=bins)
    assert check_tree(nodes, src, depth, bounds, ps.primitives)[0] == 20_000
    assert np.array_equal(bounds, ref[3])
    again = build(lib, ps.primitives, tuning=32)          # the default, spelled out: the same bytes as with no override
    assert np.array_equal(again[0], ref[0]) and np.array_equal(again[1], ref[1])


@pytest.mark.parametrize("builder", BUILDERS)
def test_optimal_slot_assignment_builds_a_valid_tree(lib, builder):
    """CRT_BVH_EXACT_SLOTS (tuning) only permutes the children of a node among its 8 slots."""
    ps = scenes.mesh_scene(64, 48, n_lon=96, n_lat=48)
    ref = build(lib, ps.primitives, tuning=builder)
    nodes, src, depth, bounds = build(lib, ps.primitives, tuning=builder | EXACT_SLOTS)
    assert check_tree(nodes, src, depth, bounds, ps.primitives)[0] == 2 * 96 * 48
    assert nodes.shape == ref[0].shape and depth == ref[2]          # same nodes, children (and leaf order) permuted
    assert not np.array_equal(nodes, ref[0])
