"""Parity of the CUDA path (through the C ABI) with the CPU oracle.  Run on the B200: -m gpu.

Bars (BASELINE.json north_star): hit indices bit-exact on fixed ray sets; accumulated radiance
max abs error <= 1e-3 and RMSE <= 1e-4 at equal spp and seed.  The numerics contract makes the
CUDA path bit-identical to the oracle, so the tests also report/require exact equality where the
design promises it.
"""
import numpy as np
import pytest

from computeraytracer_b200 import scenes
from computeraytracer_b200.renderer import (SHARD_SPLIT_SPP, SHARD_TILES, CrtError, Renderer, default_params)

pytestmark = pytest.mark.gpu

MAX_ABS_TOL = 1e-3   # north_star: per-pixel max abs error of accumulated radiance (acc / spp)
RMSE_TOL = 1e-4      # north_star: RMSE at equal spp and seed


def radiance_errors(acc_gpu, acc_cpu, spp):
    a = acc_gpu[..., :3].astype(np.float64) / spp
    b = acc_cpu[..., :3].astype(np.float64) / spp
    both_nan = np.isnan(a) & np.isnan(b)   # NaN == NaN
    d = np.where(both_nan, 0.0, a - b)
    return float(np.max(np.abs(d))), float(np.sqrt(np.mean(d * d)))


def assert_radiance(acc_gpu, acc_cpu, spp, exact=True):
    mx, rmse = radiance_errors(acc_gpu, acc_cpu, spp)
    assert mx <= MAX_ABS_TOL and rmse <= RMSE_TOL, (mx, rmse)
    if exact:
        same = np.array_equal(acc_gpu[..., :3].view(np.uint32), acc_cpu[..., :3].view(np.uint32))
        if not same:
            diff = np.argwhere(acc_gpu[..., :3].view(np.uint32) != acc_cpu[..., :3].view(np.uint32))
            pytest.fail(f"accumulators differ in {len(diff)} lanes, first at {diff[0]} (max abs {mx}, rmse {rmse})")


@pytest.fixture(scope="module")
def gpu():
    r = Renderer(0)
    yield r
    r.close()


def test_detmath_bit_exact(gpu, oracle):
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.uniform(0, 6.2831853, 100000), rng.uniform(-90, 90, 100000), rng.uniform(0.003, 1.0, 100000),
                        [0.0, -0.0, 1.0, 88.8, -87.4, -103.0, np.nan, np.inf, -np.inf]]).astype(np.float32)
    g = gpu.test_detmath(x)
    c = oracle.detmath(x)
    for a, b, name in zip(g, c, ("sin", "cos", "exp", "pow")):
        # NaN payloads differ between x86 and the GPU (0x7fc00000 vs 0x7fffffff); every consumer only
        # tests "is NaN", so NaN == NaN here and everything else is compared bit for bit
        assert np.array_equal(np.isnan(a), np.isnan(b)), name
        ok = ~np.isnan(a)
        assert np.array_equal(a[ok].view(np.uint32), b[ok].view(np.uint32)), name


def test_cornell_hits_bit_exact(gpu, oracle):
    ps = scenes.default_scene(64, 64)
    gpu.set_scene(ps)
    rays = scenes.camera_rays(ps, 200000, seed=3)
    rng = np.random.default_rng(1)
    exclude = rng.integers(0, 19, size=rays.shape[0]).astype(np.uint32)
    exclude[exclude == 18] = 0xFFFFFFFF
    sc = oracle.scene(ps)
    for ex in (None, exclude):
        gi, gt = gpu.trace_rays(rays, ex)
        ci, ct, _ = sc.trace_rays(rays, ex)
        assert np.array_equal(gi, ci)
        assert np.array_equal(gt.view(np.uint32), ct.view(np.uint32))
    assert len(np.unique(ci)) >= 17


def test_cornell_primary_histogram_matches_survey_pin(gpu, oracle):
    """the 512x512 primary-ray hit histogram, via the oracle's primary rays traced on the GPU."""
    ps = scenes.default_scene(512, 512)
    sc = oracle.scene(ps)
    idx_cpu, _ = sc.primary_hits(1)
    gpu.set_scene(ps)
    # rebuild the same primary rays on the host from the oracle probes is slow; compare through a 1-spp
    # render instead: counters must agree exactly
    gpu.reset(); gpu.reset_counters(); gpu.enable_stats(True)
    gpu.set_params(max_depth=4)
    gpu.frame(1)
    cg = gpu.counters()
    _, _, cc = oracle.scene(ps, default_params(max_depth=4)).render(1, want_fb=False)
    gpu.enable_stats(False)
    assert cg["paths"] == 512 * 512 == int(cc[0])
    assert cg["term_miss"] + cg["term_rr"] + cg["term_light"] + cg["term_depth"] == cg["paths"]
    for i, k in enumerate(["paths", "rays_ext", "rays_shadow", "prim_tests", "node_visits", "tri_tests", "rand",
                           "term_miss", "term_rr", "term_light", "term_depth", "depth_sum"]):
        assert cg[k] == int(cc[i]), k
    assert int((idx_cpu == 0xFFFFFFFF).sum()) == 25462 or abs(int((idx_cpu == 0xFFFFFFFF).sum()) - 25462) <= 12


KERNELS = [pytest.param(1, id="mega"), pytest.param(2, id="wavefront")]


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("size,spp,max_depth", [((256, 256), 4, 100), ((320, 180), 3, 100), ((101, 67), 2, 4)])
def test_cornell_radiance_parity(gpu, oracle, size, spp, max_depth, kernel):
    ps = scenes.default_scene(*size)
    gpu.set_scene(ps)
    gpu.set_params(default_params(max_depth=max_depth, kernel=kernel))
    gpu.reset()
    gpu.frame(spp)
    acc = gpu.accumulator()
    fb = gpu.framebuffer()
    assert gpu.sample == spp
    acc_c, fb_c, _ = oracle.scene(ps, default_params(max_depth=max_depth)).render(spp)
    assert_radiance(acc, acc_c, spp)
    assert int(np.max(np.abs(fb.astype(np.int32) - fb_c.astype(np.int32)))) <= 1   # rgba8: <= 1 LSB
    assert np.array_equal(fb, fb_c)
    assert (fb[..., 3] == 255).all()


@pytest.mark.parametrize("kernel", KERNELS)
def test_frame_by_frame_equals_batched(gpu, oracle, kernel):
    """n x dispatch(1) (the reference's frame loop) == dispatch(n) == chunked launches, bit for bit."""
    ps = scenes.default_scene(96, 64)
    gpu.set_scene(ps)
    gpu.set_params(default_params(kernel=kernel))
    gpu.reset()
    for _ in range(5):
        gpu.frame(1)
    a1 = gpu.accumulator(); f1 = gpu.framebuffer()
    gpu.reset(); gpu.frame(5)
    a2 = gpu.accumulator(); f2 = gpu.framebuffer()
    gpu.set_params(frames_per_launch=2)
    gpu.reset(); gpu.frame(5)
    a3 = gpu.accumulator(); f3 = gpu.framebuffer()
    gpu.set_params(frames_per_launch=0)
    assert np.array_equal(a1.view(np.uint32), a2.view(np.uint32)) and np.array_equal(f1, f2)
    assert np.array_equal(a1.view(np.uint32), a3.view(np.uint32)) and np.array_equal(f1, f3)


def test_resume_from_saved_accumulator(gpu):
    """Checkpoint/resume: the seed depends only on (x, y, sample), so restoring bindings 1 and 2 is exact."""
    ps = scenes.default_scene(64, 48)
    gpu.set_scene(ps); gpu.set_params(default_params()); gpu.reset()
    gpu.frame(3)
    saved = gpu.accumulator()
    gpu.frame(2)
    want = gpu.accumulator()
    gpu.reset()
    gpu.set_binding(1, saved)
    gpu.set_binding(2, np.array([3], dtype=np.uint32))
    gpu.frame(2)
    assert gpu.sample == 5
    assert np.array_equal(gpu.accumulator().view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("maker,kw", [(scenes.mesh_scene, dict(n_lon=64, n_lat=32)), (scenes.soup_scene, dict(n=5000))])
def test_triangle_scene_hits_bit_exact_vs_linear_scan(gpu, oracle, maker, kw):
    """BVH traversal on the GPU == the oracle's linear scan (hit index and t), with and without excludes."""
    ps = maker(64, 64, **kw)
    gpu.set_scene(ps)
    st = gpu.scene_stats()
    assert st["n_triangles"] == ps.primitives.size - 6 and st["n_linear"] == 6 and st["bvh_nodes"] > 0
    rays = scenes.camera_rays(ps, 20000, seed=5)
    rng = np.random.default_rng(2)
    exclude = rng.integers(0, ps.primitives.size, size=rays.shape[0]).astype(np.uint32)
    sc = oracle.scene(ps)
    for ex in (None, exclude):
        gi, gt = gpu.trace_rays(rays, ex)
        ci, ct, _ = sc.trace_rays(rays, ex)
        assert np.array_equal(gi, ci)
        assert np.array_equal(gt.view(np.uint32), ct.view(np.uint32))
    assert (ci >= 6).sum() > 1000


def test_tie_rule_and_coplanar_light(gpu, oracle):
    ps = scenes.default_scene(32, 32)
    tri = np.array([[[100, 100, 300], [200, 100, 300], [100, 200, 300]]] * 3, dtype=np.float32)
    tri[2, :, 2] = 555.0
    ps2 = scenes._with_triangles(ps, 6, tri)
    gpu.set_scene(ps2)
    rays = np.zeros((4, 8), dtype=np.float32)
    rays[:, 0:3] = [(120, 120, 0), (120, 120, 0), (120, 120, 400), (278, 100, 280)]
    rays[:, 4:7] = [(0, 0, 1), (0, 0, 1), (0, 0, 1), (0, 1, 0)]
    ex = np.array([0xFFFFFFFF, 7, 6, 0xFFFFFFFF], dtype=np.uint32)
    gi, gt = gpu.trace_rays(rays, ex)
    ci, ct, _ = oracle.scene(ps2).trace_rays(rays, ex)
    assert gi.tolist() == ci.tolist() == [7, 6, 8, 2]   # last: the light (2) is coplanar with the ceiling (1)
    assert np.array_equal(gt.view(np.uint32), ct.view(np.uint32))


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("maker,kw,spp", [(scenes.mesh_scene, dict(n_lon=64, n_lat=32), 3), (scenes.soup_scene, dict(n=4000), 2)])
def test_triangle_scene_radiance_parity(gpu, oracle, maker, kw, spp, kernel):
    ps = maker(128, 96, **kw)
    gpu.set_scene(ps); gpu.set_params(default_params(max_depth=8, kernel=kernel)); gpu.reset(); gpu.reset_counters()
    gpu.frame(spp)
    acc = gpu.accumulator(); fb = gpu.framebuffer()
    cg = gpu.counters()
    acc_c, fb_c, cc = oracle.scene(ps, default_params(max_depth=8)).render(spp)   # linear scan
    assert_radiance(acc, acc_c, spp)
    assert np.array_equal(fb, fb_c)
    assert (cg["paths"], cg["rays_ext"], cg["rays_shadow"]) == (int(cc[0]), int(cc[1]), int(cc[2]))


def test_million_triangle_mesh_hits(gpu, oracle):
    """BASELINE config 3 scene at full size: 64K rays against the oracle's BVH (itself pinned to its linear
    scan in tests/test_oracle_bvh.py) and 1.5K rays against the oracle's linear scan over all 1,048,582 primitives."""
    ps = scenes.mesh_scene(64, 64)
    assert ps.primitives.size == 1_048_576 + 6
    gpu.set_scene(ps)
    rays = scenes.camera_rays(ps, 65536, seed=7)
    gi, gt = gpu.trace_rays(rays)
    ci, ct, _ = oracle.scene(ps, use_bvh=True).trace_rays(rays)
    assert np.array_equal(gi, ci) and np.array_equal(gt.view(np.uint32), ct.view(np.uint32))
    li, lt, _ = oracle.scene(ps).trace_rays(rays[:1536])
    assert np.array_equal(gi[:1536], li) and np.array_equal(gt[:1536].view(np.uint32), lt.view(np.uint32))
    assert (gi >= 6).sum() > 5000


@pytest.mark.parametrize("kernel", KERNELS)
def test_virtual_ranks_tiles_are_bit_identical(gpu, oracle, kernel):
    """Tile sharding: the sum of the ranks' accumulators (each zero off its tiles) == the single-GPU result."""
    ps = scenes.default_scene(100, 52)   # ragged edge tiles on both axes
    gpu.set_scene(ps); gpu.set_params(default_params(kernel=kernel)); gpu.set_shard(0, 1, 0); gpu.reset()
    gpu.frame(3)
    want = gpu.accumulator()
    total = np.zeros_like(want)
    for world in (2, 3):
        total[:] = 0
        for r in range(world):
            gpu.set_shard(r, world, SHARD_TILES); gpu.reset(); gpu.frame(3)
            part = gpu.accumulator()
            assert np.all((part == 0) | (total == 0))
            total += part
        assert np.array_equal(total.view(np.uint32), want.view(np.uint32))
    gpu.set_shard(0, 1, 0)


@pytest.mark.parametrize("kernel", KERNELS)
def test_virtual_ranks_split_spp(gpu, oracle, kernel):
    """Split-spp: rank r renders the frames with (sample-1) % world == r; the sum matches within tolerance."""
    ps = scenes.default_scene(80, 60)
    gpu.set_scene(ps); gpu.set_params(default_params(kernel=kernel)); gpu.set_shard(0, 1, 0); gpu.reset()
    gpu.frame(6)
    want = gpu.accumulator()
    total = np.zeros_like(want, dtype=np.float32)
    for r in range(4):
        gpu.set_shard(r, 4, SHARD_SPLIT_SPP); gpu.reset(); gpu.frame(6)
        assert gpu.sample == 6
        part = gpu.accumulator()
        cpu_part, _, _ = oracle.scene(ps).render(6, rank=r, world=4, mode=SHARD_SPLIT_SPP, want_fb=False)
        assert np.array_equal(part[..., :3].view(np.uint32), cpu_part[..., :3].view(np.uint32))
        total += part
    gpu.set_shard(0, 1, 0)
    mx, rmse = radiance_errors(total, want, 6)
    assert mx <= MAX_ABS_TOL and rmse <= RMSE_TOL


def test_wavefront_detailed_counters_match_oracle(gpu, oracle):
    """Every device counter of the wavefront path on a triangle scene equals the oracle's (BVH counters aside:
    the two trees differ, only rays / tests of the linear list / rand calls / termination mix are comparable)."""
    ps = scenes.mesh_scene(96, 64, n_lon=64, n_lat=32)
    gpu.set_scene(ps); gpu.set_params(default_params(max_depth=8, kernel=2)); gpu.reset(); gpu.reset_counters()
    gpu.enable_stats(True)
    gpu.frame(2)
    cg = gpu.counters()
    gpu.enable_stats(False)
    _, _, cc = oracle.scene(ps, default_params(max_depth=8), use_bvh=True).render(2, want_fb=False)
    for i, k in enumerate(["paths", "rays_ext", "rays_shadow", "prim_tests", None, None, "rand", "term_miss", "term_rr",
                           "term_light", "term_depth", "depth_sum"]):
        if k:
            assert cg[k] == int(cc[i]), k
    assert cg["node_visits"] > 0 and cg["tri_tests"] > 0 and cg["wf_iterations"] > 0


def test_error_behaviour(gpu):
    r = Renderer(0)
    with pytest.raises(CrtError) as e:
        r.frame(1)
    assert e.value.code == 4   # CRT_ERR_STATE: bindings missing
    with pytest.raises(CrtError) as e:
        r.set_binding(4, b"\0" * 60)
    assert e.value.code == 3
    with pytest.raises(CrtError) as e:
        r.set_binding(9, b"\0" * 16)
    assert e.value.code == 2
    with pytest.raises(CrtError) as e:
        r.set_binding(8, b"\0" * 81)
    assert e.value.code == 3
    ps = scenes.default_scene(16, 16)
    r.set_scene(ps)
    r.frame(0)
    assert r.sample == 0
    with pytest.raises(CrtError):
        r.accumulator(np.empty((16, 15, 4), dtype=np.float32))
    r.close()


def test_empty_primitive_list_is_all_misses(gpu):
    ps = scenes.default_scene(24, 16)
    import dataclasses
    ps = dataclasses.replace(ps, primitives=ps.primitives[:0])
    gpu.set_scene(ps); gpu.set_params(default_params()); gpu.reset(); gpu.reset_counters()
    gpu.frame(2)
    assert not gpu.accumulator()[..., :3].any()
    c = gpu.counters()
    assert c["paths"] == 24 * 16 * 2 and c["rays_ext"] == c["paths"] and c["rays_shadow"] == 0


# ------------------------------------------------------------------------------------------------ full-size properties
# BASELINE.json's scale configs cannot be rendered by the linear-scan oracle, so at full size the CUDA path is checked
# through properties that do not depend on size: all three kernel forms agree bit for bit, the tile partition sums to
# the unsharded result, and a 1/64 sample of the image is compared with the oracle's BVH path (itself pinned to its
# linear scan in tests/test_oracle_bvh.py).

@pytest.fixture(scope="module")
def mesh1m_1080p():
    return scenes.mesh_scene(1920, 1080)


def _render(gpu, spp, **params):
    gpu.set_params(default_params(**params)); gpu.reset(); gpu.reset_counters()
    gpu.frame(spp)
    c = gpu.counters()
    return gpu.accumulator(), gpu.framebuffer(), (c["paths"], c["rays_ext"], c["rays_shadow"])


def test_full_size_mesh1m_kernel_forms_agree(gpu, mesh1m_1080p):
    """Config 3 at full size: wavefront (fused), wavefront (split kernels) and megakernel give identical bits."""
    gpu.set_scene(mesh1m_1080p); gpu.set_shard(0, 1, 0)
    a_wf, f_wf, c_wf = _render(gpu, 2, kernel=2)
    a_sp, f_sp, c_sp = _render(gpu, 2, kernel=2, flags=4)          # CRT_FLAG_WF_SPLIT
    a_mk, f_mk, c_mk = _render(gpu, 2, kernel=1)
    assert c_wf == c_sp == c_mk and c_wf[0] == 1920 * 1080 * 2
    assert np.array_equal(a_wf.view(np.uint32), a_sp.view(np.uint32)) and np.array_equal(f_wf, f_sp)
    assert np.array_equal(a_wf.view(np.uint32), a_mk.view(np.uint32)) and np.array_equal(f_wf, f_mk)
    assert a_wf[..., :3].any() and np.isfinite(a_wf[..., :3]).all()


def test_full_size_mesh1m_tiles_sum_and_oracle_sample(gpu, oracle, mesh1m_1080p):
    """Config 3 at full size: 8 tile shards sum to the unsharded accumulator; the tiles of one shard of 64 match the
    oracle's BVH render of the same tiles bit for bit."""
    gpu.set_scene(mesh1m_1080p); gpu.set_shard(0, 1, 0)
    want, _, _ = _render(gpu, 2, kernel=2)
    total = np.zeros_like(want)
    for r in range(8):
        gpu.set_shard(r, 8, SHARD_TILES)
        part, _, _ = _render(gpu, 2, kernel=2)
        assert np.all((part == 0) | (total == 0))
        total += part
    gpu.set_shard(0, 1, 0)
    assert np.array_equal(total.view(np.uint32), want.view(np.uint32))
    cpu, _, cc = oracle.scene(mesh1m_1080p, use_bvh=True).render(2, rank=5, world=64, mode=SHARD_TILES, want_fb=False)
    own = cpu[..., :3].any(axis=-1)
    assert own.sum() > 10000 and int(cc[0]) > 60000
    assert np.array_equal(want[own][:, :3].view(np.uint32), cpu[own][:, :3].view(np.uint32))


def test_soup_200k_radiance_parity_vs_oracle_bvh(gpu, oracle):
    """A mid-size random soup (config 5's generator at 200K triangles, max_depth 8) against the oracle's BVH path."""
    ps = scenes.soup_scene(160, 120, n=200_000)
    gpu.set_scene(ps); gpu.set_shard(0, 1, 0)
    acc, fb, cg = _render(gpu, 2, kernel=2, max_depth=8)
    acc_c, fb_c, cc = oracle.scene(ps, default_params(max_depth=8), use_bvh=True).render(2)
    assert_radiance(acc, acc_c, 2)
    assert np.array_equal(fb, fb_c)
    assert cg == (int(cc[0]), int(cc[1]), int(cc[2]))


@pytest.mark.parametrize("wf_slots,chunk", [(1024, 0), (4096, 2), (0, 3), (100_000, 7)])
def test_wavefront_pool_and_chunk_size_never_change_results(gpu, oracle, wf_slots, chunk):
    """The path-slot pool (how many paths are in flight, which slot runs which pixel-sample) and the chunk length (how
    many frames share a wavefront) are scheduling only: accumulator, framebuffer and counters stay those of the oracle."""
    ps = scenes.mesh_scene(96, 64, n_lon=64, n_lat=32)
    gpu.set_scene(ps); gpu.set_shard(0, 1, 0)
    acc, fb, cg = _render(gpu, 7, kernel=2, max_depth=8, wf_slots=wf_slots, frames_per_launch=chunk)
    acc_c, fb_c, cc = oracle.scene(ps, default_params(max_depth=8)).render(7)
    assert_radiance(acc, acc_c, 7)
    assert np.array_equal(fb, fb_c)
    assert cg == (int(cc[0]), int(cc[1]), int(cc[2]))


@pytest.mark.parametrize("routed,defer", [(0, 0), (1, 0), (2, 0), (2, 1), (1, 1)])
@pytest.mark.parametrize("maker,kw", [(scenes.mesh_scene, dict(n_lon=64, n_lat=32)), (scenes.default_scene, {})])
def test_wavefront_scheduling_forms_never_change_results(gpu, oracle, monkeypatch, maker, kw, routed, defer):
    """How wf_logic's live slots reach the next iteration (0: one live list; 1: finish and bounce lists filled by wf_logic
    and wf_traverse; 2: class bytes + wf_partition -- in 1 and 2 a warp's lanes are in one phase, and with `defer` a path
    that ends in a bounce warp is finished by the next iteration's finish warps) decides which thread runs which
    pixel-sample when, nothing else: accumulator, framebuffer and counters stay the oracle's.  A small pool forces many
    rounds per slot."""
    monkeypatch.setenv("CRT_WF_ROUTED", str(routed))
    monkeypatch.setenv("CRT_WF_DEFER", str(defer))
    ps = maker(96, 64, **kw)
    gpu.set_scene(ps); gpu.set_shard(0, 1, 0)
    acc, fb, cg = _render(gpu, 7, kernel=2, max_depth=8, wf_slots=2048, frames_per_launch=0)
    acc_c, fb_c, cc = oracle.scene(ps, default_params(max_depth=8)).render(7)
    assert_radiance(acc, acc_c, 7)
    assert np.array_equal(fb, fb_c)
    assert cg == (int(cc[0]), int(cc[1]), int(cc[2]))
