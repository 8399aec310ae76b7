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
    for bad in (dict(wf_tuning=4 | 8), dict(bvh_tuning=0x400 | 0x800), dict(kernel=7), dict(grid_size=0)):   # contradictory / unknown settings
        with pytest.raises(CrtError) as e:
            r.set_params(default_params(**bad))
        assert e.value.code == 1   # CRT_ERR_ARG
    with pytest.raises(CrtError) as e:
        r.combine(1)               # one rank: nothing to exchange, but a destination that does not exist is still an error
    assert e.value.code == 1
    r.combine(0)                   # ... and rank 0 of 1 is a no-op that leaves the image as it is
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


@pytest.mark.parametrize("tail", [0, 31 << 8], ids=["tail_default", "tail_off"])
@pytest.mark.parametrize("wf_slots,chunk", [(1024, 0), (4096, 2), (0, 3), (100_000, 7)])
def test_wavefront_pool_and_chunk_size_never_change_results(gpu, oracle, wf_slots, chunk, tail):
    """The path-slot pool (how many paths are in flight, which slot runs which pixel-sample) and the chunk length (how
    many frames share a wavefront) are scheduling only: accumulator, framebuffer and counters stay those of the oracle."""
    ps = scenes.mesh_scene(96, 64, n_lon=64, n_lat=32)
    gpu.set_scene(ps); gpu.set_shard(0, 1, 0)
    acc, fb, cg = _render(gpu, 7, kernel=2, max_depth=8, wf_slots=wf_slots, frames_per_launch=chunk, wf_tuning=tail)
    acc_c, fb_c, cc = oracle.scene(ps, default_params(max_depth=8)).render(7)
    assert_radiance(acc, acc_c, 7)
    assert np.array_equal(fb, fb_c)
    assert cg == (int(cc[0]), int(cc[1]), int(cc[2]))


@pytest.mark.parametrize("tail", [0, 31 << 8, 5 << 8, 13 << 8], ids=["tail_default", "tail_off", "tail_16", "tail_4096"])
@pytest.mark.parametrize("host_loop", [0, 16, 64, 32], ids=["graph_loop", "host_loop", "graph_two_lanes", "graph_one_lane"])
@pytest.mark.parametrize("live,defer", [(1, 4), (2, 4), (3, 4), (3, 8), (2, 8)])
@pytest.mark.parametrize("maker,kw", [(scenes.mesh_scene, dict(n_lon=64, n_lat=32)), (scenes.default_scene, {})])
def test_wavefront_scheduling_forms_never_change_results(gpu, oracle, maker, kw, live, defer, host_loop, tail):
    """How wf_logic's live slots reach the next iteration (crt_params.wf_tuning: 1 one live list; 2 finish and bounce lists
    filled by wf_logic and wf_traverse; 3 class bytes + wf_partition -- in 2 and 3 a warp's lanes are in one phase, and with
    CRT_WF_DEFER_ON a path that ends in a bounce warp is finished by the next iteration's finish warps) and who runs the
    iteration loop (the CUDA-graph WHILE node on the device -- over one pool, or over two half pools on two streams that
    split the chunk's samples -- or the host) and when it hands the last live slots to wf_tail
    (never, at 16 / 4096 / the default 131072 slots) decide which thread runs which pixel-sample when, nothing else:
    accumulator, framebuffer and counters stay the oracle's.  A small pool forces many rounds per slot."""
    ps = maker(96, 64, **kw)
    gpu.set_scene(ps); gpu.set_shard(0, 1, 0)
    acc, fb, cg = _render(gpu, 7, kernel=2, max_depth=8, wf_slots=2048, frames_per_launch=0, wf_tuning=live | defer | host_loop | tail)
    acc_c, fb_c, cc = oracle.scene(ps, default_params(max_depth=8)).render(7)
    assert_radiance(acc, acc_c, 7)
    assert np.array_equal(fb, fb_c)
    assert cg == (int(cc[0]), int(cc[1]), int(cc[2]))


def test_dispatch_returns_before_the_stream_drains(gpu):
    """include/crt.h: crt_dispatch_frames is asynchronous on the context's stream.  The wavefront iteration loop ends on the
    device (a CUDA-graph WHILE node), so the call only enqueues: it returns while the GPU is still rendering, and the
    work is all there after crt_synchronize.  (Round 1 read the live count back inside the loop and blocked.)"""
    import time
    ps = scenes.mesh_scene(1280, 720, n_lon=256, n_lat=128)
    gpu.set_scene(ps); gpu.set_shard(0, 1, 0)
    gpu.set_params(default_params(kernel=2)); gpu.reset(); gpu.reset_counters()
    gpu.frame(48); gpu.synchronize()                      # warm-up at the same size: pool and staging allocations, graph instantiation
    gpu.reset(); gpu.reset_counters()
    t0 = time.perf_counter(); gpu.frame(48); t_call = time.perf_counter() - t0
    t0 = time.perf_counter(); gpu.synchronize(); t_wait = time.perf_counter() - t0
    c = gpu.counters()
    assert c["paths"] == 1280 * 720 * 48 and gpu.sample == 48
    assert t_wait > 4 * t_call, (t_call, t_wait)          # the call came back long before the frames were done
    # the host-driven loop does block (that is what it is for: events between launches)
    gpu.set_params(default_params(kernel=2, wf_tuning=16)); gpu.reset()
    t0 = time.perf_counter(); gpu.frame(48); t_call_h = time.perf_counter() - t0
    gpu.synchronize()
    assert t_call_h > 10 * t_call, (t_call, t_call_h)
    gpu.set_params(default_params())


# ------------------------------------------------------------------------------------------------ builders
def _tree_ok(gpu, ps):
    """The tree that is resident on the device passes the builder invariants (tests/test_bvh_builder.py check_tree)."""
    from tests.test_bvh_builder import check_tree
    nodes, idx, info, bounds = gpu.read_bvh()
    pos_of = {int(v): k for k, v in enumerate(ps.primitives["data4"][:, 3])}
    src = np.array([pos_of[int(i)] for i in idx], dtype=np.uint32)
    pad = max(float(np.max(bounds[3:6] - bounds[0:3])), float(np.max(np.abs(bounds[0:6]))), 1.0) / 32768.0 * 0.99
    b = bounds.copy(); b[6] = pad      # (crt_read_bvh does not return the pad; this is its lower bound from the padded root box)
    return check_tree(nodes, src, int(info[2]), b, ps.primitives)


@pytest.mark.parametrize("builder", [0x400, 0x800], ids=["host_sah", "device_ploc"])
@pytest.mark.parametrize("maker,kw", [(scenes.mesh_scene, dict(n_lon=128, n_lat=64)), (scenes.soup_scene, dict(n=30_000))])
def test_builders_give_valid_trees_and_identical_hits(gpu, oracle, maker, kw, builder):
    """Host binned-SAH build and device PLOC build: the resident tree satisfies the builder invariants, and hit index and t
    equal the oracle's linear scan bit for bit (tree shape never changes a result)."""
    ps = maker(64, 64, **kw)
    gpu.set_scene(ps); gpu.set_shard(0, 1, 0); gpu.set_params(default_params(bvh_tuning=builder))
    st = gpu.scene_stats()
    assert st["builder"] == ("host_sah" if builder == 0x400 else "device_ploc") and st["bvh_nodes"] > 0
    n_tri, leaves = _tree_ok(gpu, ps)
    assert n_tri == ps.primitives.size - 6
    rays = scenes.camera_rays(ps, 20000, seed=9)
    rng = np.random.default_rng(4)
    exclude = rng.integers(0, ps.primitives.size, size=rays.shape[0]).astype(np.uint32)
    sc = oracle.scene(ps)
    for ex in (None, exclude):
        gi, gt = gpu.trace_rays(rays, ex)
        ci, ct, _ = sc.trace_rays(rays, ex)
        assert np.array_equal(gi, ci) and np.array_equal(gt.view(np.uint32), ct.view(np.uint32))
    acc, fb, cg = _render(gpu, 2, kernel=2, max_depth=8, bvh_tuning=builder)
    acc_c, fb_c, cc = oracle.scene(ps, default_params(max_depth=8)).render(2)
    assert_radiance(acc, acc_c, 2)
    assert np.array_equal(fb, fb_c) and cg == (int(cc[0]), int(cc[1]), int(cc[2]))
    gpu.set_params(default_params())


def test_host_tree_cache_is_not_fooled_by_a_reused_address(gpu, oracle):
    """The host-built tree is cached per binding-8 buffer (a group builds it once for all members).  A cache keyed by the
    buffer's ADDRESS matched a new scene allocated where a freed one had been and traversed a stale tree (found in round 2 as an
    illegal access in the repack kernel); it is keyed by the owning pointer now.  Scenes of different triangle counts set one
    after the other, many times, so that the allocator reuses addresses: hits always equal the oracle's linear scan."""
    scs = [scenes.mesh_scene(32, 32, n_lon=64, n_lat=32), scenes.soup_scene(32, 32, n=4098), scenes.mesh_scene(32, 32, n_lon=32, n_lat=64)]
    rays = scenes.camera_rays(scs[0], 3000, seed=6)
    want = [oracle.scene(ps).trace_rays(rays) for ps in scs]
    gpu.set_shard(0, 1, 0); gpu.set_params(default_params(bvh_tuning=0x400))
    for rep in range(12):
        k = rep % 3
        gpu.set_scene(scs[k])
        gi, gt = gpu.trace_rays(rays)
        assert np.array_equal(gi, want[k][0]) and np.array_equal(gt.view(np.uint32), want[k][1].view(np.uint32)), rep
    gpu.set_params(default_params())


def test_device_build_falls_back_when_the_tree_is_deeper_than_the_stack(gpu, oracle):
    """300 nested triangles around one centroid: the agglomerative build chains them deeper than the 32-entry traversal stack,
    so upload_scene takes the host SAH build (median splits bound its depth) -- and the hits are the linear scan's."""
    base = np.array([[0, 0, 0], [3, 0, 0], [0, 3, 0]], dtype=np.float32) - np.array([1, 1, 0], dtype=np.float32)
    same = np.repeat(base[None], 300, axis=0) * np.linspace(1, 2, 300, dtype=np.float32)[:, None, None] * 40.0 + np.array([278, 200, 280], dtype=np.float32)
    ps = scenes._with_triangles(scenes.default_scene(32, 32), 6, same)
    gpu.set_scene(ps); gpu.set_shard(0, 1, 0); gpu.set_params(default_params(bvh_tuning=0x800))
    st = gpu.scene_stats()
    assert st["builder"] == "host_sah" and st["bvh_depth"] + 2 <= 32
    rays = scenes.camera_rays(ps, 4000, seed=2)
    gi, gt = gpu.trace_rays(rays)
    ci, ct, _ = oracle.scene(ps).trace_rays(rays)
    assert np.array_equal(gi, ci) and np.array_equal(gt.view(np.uint32), ct.view(np.uint32))
    assert (gi >= 6).sum() > 100
    gpu.set_params(default_params())


@pytest.mark.parametrize("builder", [0x400, 0x800], ids=["host_sah", "device_ploc"])
@pytest.mark.parametrize("shift", [3.0e3, 2.0e4])
def test_mesh_far_from_the_origin_bvh_equals_linear_scan(gpu, oracle, builder, shift):
    """the conservative-culling pad was tied to the mesh's extent only; at |coordinate| / extent above ~2^9 it fell
    below one ulp of the coordinates and vanished, so rays near box faces could be culled where the linear scan accepts
    them.  A mesh of extent ~3 translated by 1e3 .. 1e4 times its extent, rays starting next to it and far from it:
    BVH == linear scan, index and t, bit for bit."""
    tris = scenes.displaced_sphere_triangles(96, 48, seed=3, center=(0.0, 0.0, 0.0), radius=1.5)
    tris = (tris.astype(np.float64) + np.array([shift, -shift * 0.5, 0.25 * shift])).astype(np.float32)
    ps = scenes._with_triangles(scenes.default_scene(32, 32), 6, tris)
    gpu.set_scene(ps); gpu.set_shard(0, 1, 0); gpu.set_params(default_params(bvh_tuning=builder))
    rng = np.random.default_rng(int(shift))
    c = np.array([shift, -shift * 0.5, 0.25 * shift])
    n = 30000
    rays = np.zeros((n, 8), dtype=np.float32)
    tgt = c + rng.normal(size=(n, 3)) * 0.9                       # aim into and around the mesh, many grazing its silhouette
    org = np.where(rng.random((n, 1)) < 0.5, c + rng.normal(size=(n, 3)) * 4.0, rng.uniform(0, 555, size=(n, 3)))
    d = tgt - org
    rays[:, 0:3] = org; rays[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    gi, gt = gpu.trace_rays(rays)
    ci, ct, _ = oracle.scene(ps).trace_rays(rays)
    assert (ci >= 6).sum() > 5000
    assert np.array_equal(gi, ci) and np.array_equal(gt.view(np.uint32), ct.view(np.uint32))
    gpu.set_params(default_params())


def test_adversarial_rays_bit_exact_vs_linear_scan(gpu, oracle):
    """SURVEY 7.2-8 / VERDICT r1: NaN components, +-0 and axis-parallel directions, origins on a patch (t == t_min region),
    |n.d| just above / below the 1e-4 cut-off (CS:546), grazing sphere rays (CS:602-612), coincident triangles and rays in
    a triangle's plane (CS:557, 566) -- hit index and t equal the oracle's linear scan bit for bit, on the reference's scene
    and on a scene with a BVH (the Wide8 slab test clamps |1/d| and skips NaN rays precisely for these)."""
    base = scenes.default_scene(32, 32)
    tri = np.array([[[100, 100, 300], [200, 100, 300], [100, 200, 300]]] * 2 + [[[150, 50, 250], [150, 250, 250], [150, 50, 450]]], dtype=np.float32)
    soup = scenes.soup_triangles(3000, seed=9)
    ps_tri = scenes._with_triangles(base, 6, np.concatenate([tri, soup]))
    rng = np.random.default_rng(12)
    rows = []

    def add(o, d):
        rows.append(list(o) + [0.0] + list(d) + [0.0])

    nan, inf = float("nan"), float("inf")
    for k in range(3):                                             # NaN / inf in every component of origin and direction
        for bad in (nan, inf, -inf):
            o = [278.0, 273.0, -800.0]; d = [0.0, 0.0, 1.0]
            o2 = list(o); o2[k] = bad; add(o2, d)
            d2 = list(d); d2[k] = bad; add(o, d2)
    for ax in range(3):                                            # axis-parallel, +0 / -0 in the other components
        for sgn in (1.0, -1.0):
            for z in (0.0, -0.0):
                for _ in range(40):
                    d = [z, z, z]; d[ax] = sgn
                    add(rng.uniform(1, 554, 3), d)
                    add([150.0, float(rng.uniform(50, 250)), float(rng.uniform(250, 450))][::1], d)   # some inside the plane x = 150 of a triangle
    for _ in range(300):                                           # origins exactly on the walls / floor / the light's plane
        p = rng.uniform(0, 555, 3); p[int(rng.integers(0, 3))] = float(rng.choice([0.0, 555.0]))
        v = rng.normal(size=3); add(p, v / np.linalg.norm(v))
    for eps in (0.99e-4, 1.0e-4, 1.01e-4, 0.5e-4, 2e-4, -0.99e-4, -1.01e-4):   # |n.d| around 0.0001 against the floor (n = +-y)
        for _ in range(20):
            phi = rng.uniform(0, 2 * np.pi); h = np.sqrt(1 - eps * eps)
            add([float(rng.uniform(100, 400)), 0.5 if eps < 0 else 0.0005, float(rng.uniform(100, 400))], [h * np.cos(phi), eps, h * np.sin(phi)])
    cs, r = np.array([190.0, 90.0, 190.0]), 90.0                   # grazing the spheres (disc ~ 0): aim at the silhouette +- ulps
    for centre, rad in ((cs, r),):
        for _ in range(300):
            o = np.array([278.0, 273.0, -800.0])
            to_c = centre - o; L = np.linalg.norm(to_c); u = to_c / L
            perp = np.cross(u, rng.normal(size=3)); perp /= np.linalg.norm(perp)
            sin_a = rad / L * (1.0 + rng.normal() * 2e-7)
            dd = u * np.sqrt(max(0.0, 1 - sin_a * sin_a)) + perp * sin_a
            add(o, dd / np.linalg.norm(dd))
    for _ in range(200):                                           # rays through the coincident pair's edges and vertices
        a, b = rng.random(2); e = rng.integers(0, 3)
        p = [(100 + 100 * a, 100, 300), (100, 100 + 100 * a, 300), (100 + 100 * a, 200 - 100 * a, 300)][int(e)]
        add([p[0], p[1], 0.0], [0.0, 0.0, 1.0])
        add([p[0], p[1], 300.0 - 50.0 * b], [0.0, 0.0, 1.0])
    rays = np.array(rows, dtype=np.float32)
    for ps in (base, ps_tri):
        gpu.set_scene(ps); gpu.set_shard(0, 1, 0)
        n_prims = ps.primitives.size
        sc = oracle.scene(ps)
        for ex in (None, rng.integers(0, min(n_prims, 12), size=rays.shape[0]).astype(np.uint32)):
            gi, gt = gpu.trace_rays(rays, ex)
            ci, ct, _ = sc.trace_rays(rays, ex)
            # (a NaN ray is accepted by the written sense of CS:557 and carries t = NaN; NaN payloads differ between x86 and
            # the GPU, every consumer only tests "is NaN": NaN == NaN here, everything else bit for bit)
            same_t = (gt.view(np.uint32) == ct.view(np.uint32)) | (np.isnan(gt) & np.isnan(ct))
            bad = np.nonzero((gi != ci) | ~same_t)[0]
            assert bad.size == 0, (bad[:5], rays[bad[:5]], gi[bad[:5]], ci[bad[:5]], gt[bad[:5]], ct[bad[:5]])
        assert len(np.unique(ci)) >= 8
    # the same rays as a render cannot be issued, but NaN-producing paths can: a 2-frame render of the triangle scene stays bit-exact
    acc, fb, cg = _render(gpu, 2, kernel=2, max_depth=8)
    acc_c, fb_c, cc = oracle.scene(ps_tri, default_params(max_depth=8)).render(2)
    assert_radiance(acc, acc_c, 2)


# ------------------------------------------------------------------------------------------------ triangle lights
def _tri_light_scene(w, h):
    ps = scenes.mesh_scene(w, h, n_lon=64, n_lat=32)
    lights = np.array([[[60, 500, 120], [180, 500, 120], [60, 500, 260]],          # a triangle light under the ceiling, facing down
                       [[400, 80, 100], [400, 80, 260], [400, 240, 100]]], dtype=np.float32)   # and one standing near the right wall
    return scenes.with_triangle_lights(ps, lights)


@pytest.mark.parametrize("tuning", [0, 16, (31 << 8), 16 | (31 << 8) | 2, 5 << 8, 64 | (5 << 8)],
                         ids=["default", "host_loop", "no_tail", "host_routed_no_tail", "tail_16", "two_lanes_tail_16"])
@pytest.mark.parametrize("kernel", KERNELS)
def test_triangle_lights_radiance_parity(gpu, oracle, kernel, tuning):
    """A light that is itself a triangle: its shadow rays need CLOSEST-hit traversal (the linear list cannot find the light)
    and the contribution can only be computed once the hit on the light is known.  Round 1 sent such scenes to the megakernel;
    now the wavefront form carries the path too (a closest-hit mode of the shadow queue, the contribution rebuilt at the next
    pass or inline in wf_tail) -- both forms against the oracle's linear scan, three lights (the reference's patch light and two
    triangles), bit for bit, through the loop on the device, the host loop, with and without wf_tail."""
    if kernel == 1 and tuning:
        pytest.skip("wf_tuning does not apply to the megakernel")
    ps = _tri_light_scene(112, 80)
    gpu.set_scene(ps); gpu.set_shard(0, 1, 0)
    st = gpu.scene_stats()
    assert st["shadow_closest"]
    acc, fb, cg = _render(gpu, 4, kernel=kernel, max_depth=8, wf_slots=4096 if tuning else 0, wf_tuning=tuning)
    acc_c, fb_c, cc = oracle.scene(ps, default_params(max_depth=8)).render(4)
    assert_radiance(acc, acc_c, 4)
    assert np.array_equal(fb, fb_c)
    assert cg == (int(cc[0]), int(cc[1]), int(cc[2]))
    # the triangle lights do light the scene: more radiance than with the patch light alone
    ps0 = scenes.mesh_scene(112, 80, n_lon=64, n_lat=32)
    acc0, _, _ = oracle.scene(ps0, default_params(max_depth=8)).render(4)
    assert float(np.nansum(acc_c[..., 1])) != float(np.nansum(acc0[..., 1]))


def test_triangle_lights_auto_uses_the_wavefront_form(gpu, oracle):
    ps = _tri_light_scene(96, 64)
    gpu.set_scene(ps); gpu.set_shard(0, 1, 0)
    acc, fb, cg = _render(gpu, 3, max_depth=8)                   # AUTO
    c = gpu.counters()
    assert c["wf_iterations"] > 0                               # not the megakernel
    acc_c, fb_c, cc = oracle.scene(ps, default_params(max_depth=8)).render(3)
    assert_radiance(acc, acc_c, 3)
    assert np.array_equal(fb, fb_c)


# ------------------------------------------------------------------------------------------------ single-process group
@pytest.mark.parametrize("mode", [SHARD_TILES, SHARD_SPLIT_SPP], ids=["tiles", "split_spp"])
@pytest.mark.parametrize("n", [2, 3])
def test_group_on_one_device_combines_like_one_gpu(gpu, oracle, mode, n):
    """crt_create_multi with the same device n times (a group's logic does not care where its members live): bindings are
    replicated, frames sharded, crt_read_binding combines on member 0 with the fused peer kernel.  Tiles: bit-identical to
    one context; split-spp: within the tolerance; frames / read / frames / read equals one uninterrupted render (the
    combine is repeatable: partial sums stay partial)."""
    ps = scenes.mesh_scene(100, 52, n_lon=64, n_lat=32)            # ragged edge tiles on both axes
    gpu.set_scene(ps); gpu.set_shard(0, 1, 0)
    want5, fb5, c5 = _render(gpu, 5, max_depth=8)
    gpu.frame(4)
    want9, fb9 = gpu.accumulator(), gpu.framebuffer()
    with Renderer(devices=[0] * n, shard_mode=mode) as g:
        g.set_scene(ps); g.set_params(max_depth=8); g.reset(); g.reset_counters()
        g.frame(5)
        a5, f5 = g.accumulator(), g.framebuffer()
        c = g.counters()
        assert (c["paths"], c["rays_ext"], c["rays_shadow"]) == c5 and g.sample == 5
        g.frame(4)
        a9, f9 = g.accumulator(), g.framebuffer()
        a9b = g.accumulator()                                       # reading twice combines twice: same image
        ms, nbytes = g.last_combine()
        assert np.array_equal(a9.view(np.uint32), a9b.view(np.uint32)) and g.sample == 9
        if mode == SHARD_TILES:
            assert np.array_equal(a5[..., :3].view(np.uint32), want5[..., :3].view(np.uint32)) and np.array_equal(f5, fb5)
            assert np.array_equal(a9[..., :3].view(np.uint32), want9[..., :3].view(np.uint32)) and np.array_equal(f9, fb9)
            assert 0 < nbytes < 100 * 52 * 16                       # the others' tiles only
        else:
            for got, want, spp in ((a5, want5, 5), (a9, want9, 9)):
                mx, rmse = radiance_errors(got, want, spp)
                assert mx <= MAX_ABS_TOL and rmse <= RMSE_TOL
            assert int(np.abs(f9.astype(int) - fb9.astype(int)).max()) <= 1
            assert nbytes == (n - 1) * 100 * 52 * 16
        # resume on a group: the saved image goes back in, more frames, same as uninterrupted (tiles: bit-identical)
        g.reset(); g.set_binding(1, a5); g.set_binding(2, np.array([5], dtype=np.uint32))
        g.frame(4)
        r9 = g.accumulator()
        if mode == SHARD_TILES:
            assert np.array_equal(r9[..., :3].view(np.uint32), want9[..., :3].view(np.uint32))
        else:
            mx, rmse = radiance_errors(r9, want9, 9)
            assert mx <= MAX_ABS_TOL and rmse <= RMSE_TOL


# ------------------------------------------------------------------------------------------------ two physical GPUs
def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _nccl_worker(rank, world, mode, port, out_dir):
    import os as _os
    import sys as _sys
    _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
    _os.environ["MASTER_ADDR"] = "127.0.0.1"; _os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    from computeraytracer_b200 import multi
    dist.init_process_group("gloo", rank=rank, world_size=world)       # carries the 128-byte id only; the data path is libcrt's NCCL
    ps = scenes.mesh_scene(100, 52, n_lon=64, n_lat=32)
    with Renderer(rank) as r:
        r.set_scene(ps); r.set_params(max_depth=8)
        multi.init(r, mode)
        r.reset(); r.frame(5); r.combine(0)
        if rank == 0:
            np.save(_os.path.join(out_dir, "a5.npy"), r.accumulator()); np.save(_os.path.join(out_dir, "f5.npy"), r.framebuffer())
        r.frame(4); r.combine(0); r.combine(0)                         # repeatable
        ms, nbytes = r.last_combine()
        if rank == 0:
            np.save(_os.path.join(out_dir, "a9.npy"), r.accumulator()); np.save(_os.path.join(out_dir, "f9.npy"), r.framebuffer())
        else:
            np.save(_os.path.join(out_dir, f"sent_{rank}.npy"), np.array([nbytes], dtype=np.int64))
        r.synchronize()
        dist.barrier()
        r.comm_destroy()
    dist.destroy_process_group()


@pytest.mark.skipif(_n_gpus() < 2, reason="needs two physical GPUs (gpurun --gpus 2)")
@pytest.mark.parametrize("mode", [SHARD_TILES, SHARD_SPLIT_SPP], ids=["tiles", "split_spp"])
def test_two_gpus_nccl_combine_behind_the_c_abi(gpu, tmp_path, mode):
    """One process per GPU, crt_comm_init + crt_combine (NCCL inside libcrt): tiles gathered from two physical devices are
    bit-identical to one GPU and only the owned tiles travel; split-spp reduces within the tolerance; repeatable."""
    import os
    import torch.multiprocessing as mp
    ps = scenes.mesh_scene(100, 52, n_lon=64, n_lat=32)
    gpu.set_scene(ps); gpu.set_shard(0, 1, 0)
    want5, fb5, _ = _render(gpu, 5, max_depth=8)
    gpu.frame(4)
    want9, fb9 = gpu.accumulator(), gpu.framebuffer()
    mp.spawn(_nccl_worker, args=(2, mode, 29700 + os.getpid() % 1000 + mode, str(tmp_path)), nprocs=2, join=True)
    a5, f5, a9, f9 = (np.load(tmp_path / f"{k}.npy") for k in ("a5", "f5", "a9", "f9"))
    sent = int(np.load(tmp_path / "sent_1.npy")[0])
    if mode == SHARD_TILES:
        assert np.array_equal(a5[..., :3].view(np.uint32), want5[..., :3].view(np.uint32)) and np.array_equal(f5, fb5)
        assert np.array_equal(a9[..., :3].view(np.uint32), want9[..., :3].view(np.uint32)) and np.array_equal(f9, fb9)
        assert sent == ((13 * 13) // 2) * 32 * 16                   # rank 1 owns 84 of the 169 tiles: 84 x 32 x 16 bytes
    else:
        for got, want, spp in ((a5, want5, 5), (a9, want9, 9)):
            mx, rmse = radiance_errors(got, want, spp)
            assert mx <= MAX_ABS_TOL and rmse <= RMSE_TOL
        assert sent == 100 * 52 * 16


@pytest.mark.skipif(_n_gpus() < 2, reason="needs two physical GPUs (gpurun --gpus 2)")
@pytest.mark.parametrize("mode", [SHARD_TILES, SHARD_SPLIT_SPP], ids=["tiles", "split_spp"])
def test_two_gpus_single_process_group(gpu, mode):
    """crt_create_multi over two physical devices: the fused combine reads device 1's accumulator through a peer pointer."""
    ps = scenes.mesh_scene(100, 52, n_lon=64, n_lat=32)
    gpu.set_scene(ps); gpu.set_shard(0, 1, 0)
    want, fb, c1 = _render(gpu, 6, max_depth=8)
    with Renderer(devices=[0, 1], shard_mode=mode) as g:
        g.set_scene(ps); g.set_params(max_depth=8); g.reset(); g.reset_counters()
        g.frame(6)
        a, f = g.accumulator(), g.framebuffer()
        c = g.counters()
        assert (c["paths"], c["rays_ext"], c["rays_shadow"]) == c1
        if mode == SHARD_TILES:
            assert np.array_equal(a[..., :3].view(np.uint32), want[..., :3].view(np.uint32)) and np.array_equal(f, fb)
        else:
            mx, rmse = radiance_errors(a, want, 6)
            assert mx <= MAX_ABS_TOL and rmse <= RMSE_TOL


# ------------------------------------------------------------------------------------------------ the other configs at full size
def test_full_size_config2_cornell_1080p_64spp_tile_sample(gpu, oracle):
    """BASELINE config 2 at its stated size (reference default scene, 1920x1080, 64 spp, max_depth 100): the tiles of one
    shard of 64 are bit-identical to the oracle's render of the same tiles, the counters add up."""
    ps = scenes.default_scene(1920, 1080)
    gpu.set_scene(ps); gpu.set_shard(0, 1, 0)
    acc, fb, cg = _render(gpu, 64)
    assert cg[0] == 1920 * 1080 * 64 and np.isfinite(acc[..., :3]).all()
    cpu, fb_c, cc = oracle.scene(ps).render(64, rank=11, world=64, mode=SHARD_TILES)
    own = cpu[..., :3].any(axis=-1)
    assert own.sum() > 15000
    assert np.array_equal(acc[own][:, :3].view(np.uint32), cpu[own][:, :3].view(np.uint32))
    assert np.array_equal(fb[own], fb_c[own])


def test_full_size_config4_mesh1m_4k_tile_sample(gpu, oracle):
    """BASELINE config 4's image size (3840x2160) on the 1M-triangle scene: 1/256 of the tiles against the oracle's BVH path."""
    ps = scenes.mesh_scene(3840, 2160)
    gpu.set_scene(ps); gpu.set_shard(0, 1, 0)
    acc, fb, cg = _render(gpu, 2, kernel=2)
    assert cg[0] == 3840 * 2160 * 2
    cpu, _, cc = oracle.scene(ps, use_bvh=True).render(2, rank=77, world=256, mode=SHARD_TILES, want_fb=False)
    own = cpu[..., :3].any(axis=-1)
    assert own.sum() > 10000
    assert np.array_equal(acc[own][:, :3].view(np.uint32), cpu[own][:, :3].view(np.uint32))


def test_full_size_config5_soup10m_tile_sample(gpu, oracle):
    """BASELINE config 5 at its stated size: 10,000,000 random triangles, max_depth 8, 1080p.  1/128 of the tiles against
    the oracle's BVH path (itself pinned to its linear scan in tests/test_oracle_bvh.py), and 20K rays' hits."""
    ps = scenes.soup_scene(1920, 1080)
    assert ps.primitives.size == 10_000_006
    gpu.set_scene(ps); gpu.set_shard(0, 1, 0)
    acc, fb, cg = _render(gpu, 2, kernel=2, max_depth=8)
    assert cg[0] == 1920 * 1080 * 2 and np.isfinite(acc[..., :3]).all()
    sc = oracle.scene(ps, default_params(max_depth=8), use_bvh=True)
    cpu, _, cc = sc.render(2, rank=5, world=128, mode=SHARD_TILES, want_fb=False)
    # (ten million triangles fill the box: the light is occluded almost everywhere and most pixels are exactly black, so the
    # shard's pixels are taken from the tile ownership, not from "non-zero")
    from computeraytracer_b200 import multi
    units = multi.tile_units(1920, 1080, 5, 128)
    own = np.zeros(1920 * 1080, dtype=bool); own[units[units >= 0]] = True
    own = own.reshape(1080, 1920)
    assert own.sum() > 16000 and int(cc[0]) == 2 * int(own.sum()) and int(cc[1]) > 3 * int(own.sum())
    assert not cpu[~own][:, :3].any()
    assert np.array_equal(acc[own][:, :3].view(np.uint32), cpu[own][:, :3].view(np.uint32))
    assert acc[..., :3].any()
    rays = scenes.camera_rays(ps, 20000, seed=3)
    gi, gt = gpu.trace_rays(rays)
    ci, ct, _ = sc.trace_rays(rays)
    assert np.array_equal(gi, ci) and np.array_equal(gt.view(np.uint32), ct.view(np.uint32))
