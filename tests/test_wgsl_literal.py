"""The C oracle against a second, independent reading of the shader (tests/wgsl_literal.py): bit for bit.

The reference has no golden vectors and its WGSL cannot run on this box, so the oracle is "unpinned by the reference"
(DESIGN.md section 2).  What CAN be checked is that two restatements of the shader written separately -- scalar Python that
keeps the WGSL's structure, and the C the kernels are tested against -- agree on every bit of every pixel-sample, on the
framebuffer bytes and on every statistic (rays, rand() calls, primitive tests, how each path ended): a misreading of the shader
would have to be made twice.  The shared part is stated in wgsl_literal's header: where DESIGN section 3 puts an fmaf, and the
sin / cos / exp / pow of include/crt_detmath.h.  The last test drops the fmaf choice and shows it is only rounding noise."""
import dataclasses

import numpy as np
import pytest

from computeraytracer_b200 import scenes
from computeraytracer_b200.scene_loader import PRIM_DTYPE, copy_records
from tests import orc
from tests import wgsl_literal as wl


@pytest.fixture(scope="module")
def machine_parts(oracle):
    def sincos(x):
        s, c, _, _ = oracle.detmath(np.array([x], dtype=np.float32))
        return s[0], c[0]

    def exp(x):
        return oracle.detmath(np.array([x], dtype=np.float32))[2][0]

    def pow_1_2p4(x):
        return oracle.detmath(np.array([x], dtype=np.float32))[3][0]
    return sincos, exp, pow_1_2p4


def _statistics(sh):
    """In the order of include/crt.h's CRT_CNT_* 0..11 (4 and 5 are the BVH's: zero in a scene without triangles)."""
    return [sh.paths, sh.rays_ext, sh.rays_shadow, sh.prim_tests, 0, 0, sh.n_rand, sh.term["miss"], sh.term["rr"], sh.term["light"],
            sh.term["depth"], sh.depth_sum]


def test_oracle_equals_the_literal_reading_bit_for_bit(oracle, machine_parts):
    W = H = 32
    ps = scenes.default_scene(W, H)
    sh = wl.Shader(ps.bindings(), wl.Machine(*machine_parts, fused=True))
    sc = oracle.scene(ps)
    acc_l = np.zeros((H, W, 4), dtype=np.float32)                     # alt_color_buffer (CS:2): vec3 in 16-byte slots
    for sample in (1, 2, 3):                                          # three dispatches, UpdateVariables.wgsl's sample++ before each
        for y in range(H):
            for x in range(W):
                acc_l[y, x, :3] += np.array(sh.main(x, y, sample), dtype=np.float32)
    assert sh.main(W, 0, 1) is None and sh.main(0, H, 1) is None      # CS:88: invocations outside the screen return
    acc, fb, cnt = sc.render(3, threads=1)
    assert np.array_equal(acc_l[..., :3].view(np.uint32), acc[..., :3].view(np.uint32))
    assert _statistics(sh) == [int(c) for c in cnt[:12]]
    fb_l = np.array([[sh.resolve(acc_l[y, x, :3], 3) for x in range(W)] for y in range(H)], dtype=np.uint8)
    assert np.array_equal(fb_l, fb)
    # the comparison means something only if the paths went everywhere the shader has code
    assert sh.events["reflect"] > 50 and sh.events["refract"] > 200 and sh.events["attenuate"] > 100
    assert sh.events["mis_light_hit"] > 5 and sh.term["light"] > sh.events["mis_light_hit"]
    assert sh.term["miss"] > 500 and sh.term["rr"] > 500
    assert np.count_nonzero(fb_l[..., :3]) > W * H


def test_oracle_equals_the_literal_reading_other_sample_cells_and_depth_limit(oracle, machine_parts):
    """sample = 18: `sample % GRID_SIZE` (CS:497) is 2, not the sample; MAXDEPTH = 2: the depth exit (CS:167) is taken often."""
    W, H = 24, 16
    ps = scenes.default_scene(W, H)
    sh = wl.Shader(ps.bindings(), wl.Machine(*machine_parts, fused=True), max_depth=2)
    sc = oracle.scene(ps, orc.default_params(max_depth=2))
    acc, _, cnt = sc.render(1, sample_start=17, threads=1, want_fb=False)
    got = np.array([[sh.main(x, y, 18) for x in range(W)] for y in range(H)], dtype=np.float32)
    assert np.array_equal(got.view(np.uint32), acc[..., :3].view(np.uint32))
    assert _statistics(sh) == [int(c) for c in cnt[:12]]
    assert sh.term["depth"] > 20


def test_oracle_equals_the_literal_reading_other_parameters(oracle, machine_parts):
    """crt_params that stand for the shader's constants: GRID_SIZE 4 (CS:15), t_min 0.01 (CS:720), exposure 1.0 (CS:112)."""
    W, H = 20, 20
    ps = scenes.default_scene(W, H)
    sh = wl.Shader(ps.bindings(), wl.Machine(*machine_parts, fused=True), grid_size=4)
    sc = oracle.scene(ps, orc.default_params(grid_size=4, t_min=0.01, exposure=1.0))
    acc_l = np.zeros((H, W, 3), dtype=np.float32)
    for sample in (5, 6):
        for y in range(H):
            for x in range(W):
                acc_l[y, x] += np.array(sh.main(x, y, sample, t_min=0.01), dtype=np.float32)
    acc, fb, cnt = sc.render(2, sample_start=4, threads=1)
    assert np.array_equal(acc_l.view(np.uint32), acc[..., :3].view(np.uint32))
    assert _statistics(sh) == [int(c) for c in cnt[:12]]
    fb_l = np.array([[sh.resolve(acc_l[y, x], 6, exposure=1.0) for x in range(W)] for y in range(H)], dtype=np.uint8)
    assert np.array_equal(fb_l, fb)


def _two_light_scene(W, H):
    """The default scene plus a second light whose emission index (1) is in range for `lights[emission_index]` (CS:362 looks a
    light up by its SPECTRUM index: out of range -- clamped -- for the default light, the wrong light's area for this one; both
    readings must make the same 'mistake'), a tilted glass pane (glass on a patch, both faces), a small sphere, another camera."""
    ps = scenes.default_scene(W, H)
    n = ps.primitives.shape[0]
    extra = np.zeros(3, dtype=PRIM_DTYPE)
    extra[0] = (0, (1.0, 200.0, 150.0), (0.0, 120.0, 0.0), (0.0, 0.0, 160.0), (1, 0, 1, n))
    extra[1] = (0, (120.0, 40.0, 30.0), (300.0, 40.0, 35.0), (-20.0, 260.0, 60.0), (4, 0, 2, n + 1))
    extra[2] = (1, (430.0, 40.0, 120.0), (40.0, 40.0, 40.0), (0.0, 0.0, 0.0), (4, 1, 0, n + 2))

    def raw(a):
        return np.ascontiguousarray(a).view(np.uint8).tobytes()
    cam = ps.camera.copy()
    cam[0:3] = (300.0, 320.0, -700.0)
    cam[4:7] = (260.0, 250.0, 0.0)
    cam[8:11] = (0.1, 1.0, 0.05)
    return dataclasses.replace(ps, primitives=copy_records(raw(ps.primitives) + raw(extra), PRIM_DTYPE),
                               lights=copy_records(raw(ps.lights) + raw(extra[:1]), PRIM_DTYPE), camera=cam)


def test_oracle_equals_the_literal_reading_two_lights_glass_pane(oracle, machine_parts):
    W = H = 24
    ps = _two_light_scene(W, H)
    sh = wl.Shader(ps.bindings(), wl.Machine(*machine_parts, fused=True))
    acc_l = np.zeros((H, W, 3), dtype=np.float32)
    for sample in (1, 2):
        for y in range(H):
            for x in range(W):
                acc_l[y, x] += np.array(sh.main(x, y, sample), dtype=np.float32)
    acc, fb, cnt = oracle.scene(ps).render(2, threads=1)
    assert np.array_equal(acc_l.view(np.uint32), acc[..., :3].view(np.uint32))
    assert _statistics(sh) == [int(c) for c in cnt[:12]]
    fb_l = np.array([[sh.resolve(acc_l[y, x], 2) for x in range(W)] for y in range(H)], dtype=np.uint8)
    assert np.array_equal(fb_l, fb)
    assert sh.events["refract"] > 300 and sh.events["mis_light_hit"] > 10 and sh.term["light"] > 30


def _random_scene(seed, W, H):
    """6-12 random patches and spheres with random materials and spectrum indices, 1-4 lights of either kind (a sphere as a light
    has a zero `data3`: area 0, pdf inf, NaN radiance -- as undefined in the shader as here, but the same bits everywhere), a
    random camera.  Nothing a scene file of the reference could not say."""
    rng = np.random.default_rng(seed)
    ps = scenes.default_scene(W, H)
    n = int(rng.integers(6, 13))
    prims = np.zeros(n, dtype=PRIM_DTYPE)
    for i in range(n):
        mat = 1 if i == 0 else int(rng.choice([0, 0, 0, 1, 2]))
        idx = (int(rng.integers(0, 7)), int(rng.integers(0, 7)), mat, i)
        if rng.random() < 0.6:
            prims[i] = (0, tuple(rng.uniform(0, 555, 3)), tuple(rng.uniform(-300, 300, 3)), tuple(rng.uniform(-300, 300, 3)), idx)
        else:
            r = float(rng.uniform(20, 120))
            prims[i] = (1, tuple(rng.uniform(50, 500, 3)), (r, r, r), (0, 0, 0), idx)

    def raw(a):
        return np.ascontiguousarray(a).view(np.uint8).tobytes()
    lights = prims[[i for i in range(n) if prims[i]["data4"][2] == 1]]
    cam = ps.camera.copy()
    cam[0:3] = rng.uniform(-600, 1100, 3)
    cam[4:7] = rng.uniform(150, 400, 3)
    cam[8:11] = (0, 1, 0)
    return dataclasses.replace(ps, primitives=copy_records(raw(prims), PRIM_DTYPE), lights=copy_records(raw(lights), PRIM_DTYPE), camera=cam)


def _same_bits_or_both_nan(a, b):
    return bool(((a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))).all())


@pytest.mark.parametrize("seed", [2, 3, 4, 7])
def test_oracle_equals_the_literal_reading_random_scenes(oracle, machine_parts, seed):
    W = H = 16
    ps = _random_scene(seed, W, H)
    sh = wl.Shader(ps.bindings(), wl.Machine(*machine_parts, fused=True))
    acc_l = np.zeros((H, W, 3), dtype=np.float32)
    for sample in (1, 2):
        for y in range(H):
            for x in range(W):
                acc_l[y, x] += np.array(sh.main(x, y, sample), dtype=np.float32)
    acc, fb, cnt = oracle.scene(ps).render(2, threads=1)
    assert _same_bits_or_both_nan(acc_l, acc[..., :3])
    assert _statistics(sh) == [int(c) for c in cnt[:12]]
    fb_l = np.array([[sh.resolve(acc_l[y, x], 2) for x in range(W)] for y in range(H)], dtype=np.uint8)
    assert np.array_equal(fb_l, fb)


def test_the_fmaf_choice_is_rounding_noise(oracle, machine_parts):
    """WGSL lets an implementation fuse a*b+c or not.  With every fmaf of section 3 replaced by two roundings the literal
    reading stays within 1e-4 relative of the oracle on all but a handful of pixel-samples -- those whose path took a different
    branch somewhere (a silhouette, a Russian-roulette draw): noise of the kind any two WGSL back ends show, not semantics."""
    W = H = 32
    ps = scenes.default_scene(W, H)
    sh = wl.Shader(ps.bindings(), wl.Machine(*machine_parts, fused=False))
    acc, _, _ = oracle.scene(ps).render(1, threads=1, want_fb=False)
    got = np.array([[sh.main(x, y, 1) for x in range(W)] for y in range(H)], dtype=np.float32)
    same = np.array_equal(got.view(np.uint32), acc[..., :3].view(np.uint32))
    close = np.isclose(got, acc[..., :3], rtol=1e-4, atol=1e-6).all(axis=-1)
    assert not same                                                    # the switch does something
    assert (~close).sum() <= 8, int((~close).sum())


@pytest.mark.gpu
@pytest.mark.parametrize("scene", ["default", "two_lights"])
@pytest.mark.parametrize("kernel", ["wavefront", "mega"])
def test_cuda_path_equals_the_literal_reading_bit_for_bit(oracle, machine_parts, kernel, scene):
    """The same comparison without the C oracle in the middle: libcrt through the C ABI against the Python reading."""
    from computeraytracer_b200.renderer import KERNEL_MEGA, KERNEL_WAVEFRONT, Renderer
    W = H = 20
    ps = scenes.default_scene(W, H) if scene == "default" else _two_light_scene(W, H)
    sh = wl.Shader(ps.bindings(), wl.Machine(*machine_parts, fused=True))
    acc_l = np.zeros((H, W, 3), dtype=np.float32)
    for sample in (1, 2):
        for y in range(H):
            for x in range(W):
                acc_l[y, x] += np.array(sh.main(x, y, sample), dtype=np.float32)
    fb_l = np.array([[sh.resolve(acc_l[y, x], 2) for x in range(W)] for y in range(H)], dtype=np.uint8)
    with Renderer(0) as r:
        r.set_scene(ps)
        r.set_params(kernel=KERNEL_MEGA if kernel == "mega" else KERNEL_WAVEFRONT)
        r.frame(2)
        acc, fb, c = r.accumulator(), r.framebuffer(), r.counters()
    assert np.array_equal(acc[..., :3].view(np.uint32), acc_l.view(np.uint32))
    assert np.array_equal(fb, fb_l)
    assert (c["paths"], c["rays_ext"], c["rays_shadow"]) == (sh.paths, sh.rays_ext, sh.rays_shadow)


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", ["wavefront", "mega"])
def test_cuda_path_two_lights_glass_pane_larger(oracle, kernel):
    """The variant scene at a size the Python reading would take minutes for: libcrt against the C oracle, 96x96 x 5 frames."""
    from computeraytracer_b200.renderer import KERNEL_MEGA, KERNEL_WAVEFRONT, Renderer
    ps = _two_light_scene(96, 96)
    acc_c, fb_c, cc = oracle.scene(ps).render(5)
    with Renderer(0) as r:
        r.set_scene(ps)
        r.set_params(kernel=KERNEL_MEGA if kernel == "mega" else KERNEL_WAVEFRONT)
        r.frame(2)
        r.frame(3)
        acc, fb, c = r.accumulator(), r.framebuffer(), r.counters()
    assert np.array_equal(acc[..., :3].view(np.uint32), acc_c[..., :3].view(np.uint32))
    assert np.array_equal(fb, fb_c)
    assert (c["paths"], c["rays_ext"], c["rays_shadow"]) == (int(cc[0]), int(cc[1]), int(cc[2]))


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", ["wavefront", "mega"])
def test_cuda_path_random_scenes(oracle, kernel):
    """libcrt against the C oracle on the random scenes (several lights, sphere lights, NaN radiance): accumulators bit for bit
    (NaN = NaN: the GPU's canonical NaN has another payload than the host's), framebuffer bytes, ray counts."""
    from computeraytracer_b200.renderer import KERNEL_MEGA, KERNEL_WAVEFRONT, Renderer
    W = H = 48
    with Renderer(0) as r:
        r.set_params(kernel=KERNEL_MEGA if kernel == "mega" else KERNEL_WAVEFRONT)
        for seed in range(12):
            ps = _random_scene(seed, W, H)
            acc_c, fb_c, cc = oracle.scene(ps).render(3)
            r.set_scene(ps)
            r.reset()
            r.reset_counters()
            r.frame(1)
            r.frame(2)
            acc, fb, c = r.accumulator(), r.framebuffer(), r.counters()
            assert _same_bits_or_both_nan(acc[..., :3], acc_c[..., :3]), seed
            assert np.array_equal(fb, fb_c), seed
            assert (c["paths"], c["rays_ext"], c["rays_shadow"]) == (int(cc[0]), int(cc[1]), int(cc[2])), seed


def _random_scene_with_triangles(seed, W, H):
    """A random scene (above) plus 40-400 random triangles (this project's category 2) of random materials -- diffuse, light,
    glass -- every light-material primitive also listed in the lights binding."""
    ps = _random_scene(seed, W, H)
    rng = np.random.default_rng(1000 + seed)
    n0 = ps.primitives.shape[0]
    nt = int(rng.integers(40, 401))
    tris = np.zeros(nt, dtype=PRIM_DTYPE)
    for i in range(nt):
        mat = int(rng.choice([0, 0, 0, 0, 1, 2, 2]))
        size = float(rng.choice([15.0, 60.0, 250.0]))
        tris[i] = (2, tuple(rng.uniform(0, 555, 3)), tuple(rng.uniform(-size, size, 3)), tuple(rng.uniform(-size, size, 3)),
                   (int(rng.integers(0, 7)), int(rng.integers(0, 7)), mat, n0 + i))

    def raw(a):
        return np.ascontiguousarray(a).view(np.uint8).tobytes()
    prims = copy_records(raw(ps.primitives) + raw(tris), PRIM_DTYPE)
    lights = prims[[i for i in range(prims.shape[0]) if prims[i]["data4"][2] == 1]]
    return dataclasses.replace(ps, primitives=prims, lights=copy_records(raw(lights), PRIM_DTYPE))


@pytest.mark.gpu
@pytest.mark.parametrize("builder", ["host", "device"])
@pytest.mark.parametrize("kernel", ["wavefront", "mega"])
def test_cuda_path_random_scenes_with_triangles(oracle, kernel, builder):
    """The same with triangles in the scene (BVH traversal, triangle lights, glass triangles): libcrt against the oracle's linear scan."""
    from computeraytracer_b200.renderer import BVH_BUILDER_DEVICE, BVH_BUILDER_HOST_SAH, KERNEL_MEGA, KERNEL_WAVEFRONT, Renderer
    W = H = 40
    with Renderer(0) as r:
        r.set_params(kernel=KERNEL_MEGA if kernel == "mega" else KERNEL_WAVEFRONT, max_depth=12,
                     bvh_tuning=BVH_BUILDER_DEVICE if builder == "device" else BVH_BUILDER_HOST_SAH)
        for seed in range(8):
            ps = _random_scene_with_triangles(seed, W, H)
            acc_c, fb_c, cc = oracle.scene(ps, orc.default_params(max_depth=12)).render(3)
            r.set_scene(ps)
            r.reset()
            r.reset_counters()
            r.frame(1)
            r.frame(2)
            acc, fb, c = r.accumulator(), r.framebuffer(), r.counters()
            assert _same_bits_or_both_nan(acc[..., :3], acc_c[..., :3]), seed
            assert np.array_equal(fb, fb_c), seed
            assert (c["paths"], c["rays_ext"], c["rays_shadow"]) == (int(cc[0]), int(cc[1]), int(cc[2])), seed
