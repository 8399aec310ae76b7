"""Known-answer tests pinning the CPU oracle.

The reference has no tests or golden vectors of its own (package.json:9), so these pins are
survey-derived from the cited shader lines: B.1 is exact integer arithmetic (authoritative),
B.2 exact byte layouts, B.3 f32 primary rays (exact up to the rounding pins of DESIGN.md).
"""
import hashlib

import numpy as np
import pytest

from computeraytracer_b200.scenes import default_scene


# ---- B.1 RNG, ComputeShader.wgsl:865-897, seed CS:98
TEA = [((0, 0), 0x741C187D), ((1, 0), 0x8DA6B311), ((0, 100), 0xFBA60970), ((1, 100), 0x86664F3A),
       ((500, 50000), 0xD61E8C8E), ((999, 99900), 0x3BBB9544), ((1919, 107900), 0xC0F8545B)]


@pytest.mark.parametrize("args,want", TEA)
def test_tea(oracle, args, want):
    assert oracle.tea(*args) == want


STREAMS = [
    ((0, 0, 1), [3982936, 1279938, 10369472, 15862708], [0x80F20BB4, 0x23662435, 0x85F38847, 0x07A86206]),
    ((1, 0, 1), [14317009, 12399345, 13684972, 15672310], [0xECEF23F6, 0xBD4A81DC, 0xD1C7B785, 0xF5F00255]),
    ((0, 1, 1), [15065161, 7315900, 9490430, 9311275], [0xC88E142B, 0xE6767C63, 0x47B60546, 0x90541108]),
    ((500, 500, 1), [7913910, 12317604, 9026109, 599401], [0xEC092569, 0xD4459978, 0xF0E0D0F8, 0x4D0FC440]),
    ((500, 500, 2), [6150118, 1728512, 11137723, 11349074], [0x0FAD2C52, 0x6619A8BC, 0x809F3D8A, 0xE761A163]),
    ((999, 999, 1), [1147759, 7560048, 8088559, 11396148], [0xC5ADE434, 0x7054D159, 0x49BB2E1F, 0x34781327]),
    ((1919, 1079, 64), [1288756, 5462617, 9844093, 10010033], [0xEA98BDB1, 0x3FD5AA3A, 0x9A7F947D, 0xA5381EFC]),
]


@pytest.mark.parametrize("pxs,first4,state", STREAMS)
def test_rand_stream(oracle, pxs, first4, state):
    got, st = oracle.rand_stream(*pxs, 4)
    assert got == first4
    assert st == state


def test_seed_layout(oracle):
    s = oracle.seed(0, 0, 1)
    assert list(s) == [0, 0, 1, 0x741C187D]
    s = oracle.seed(7, 3, 9)
    assert list(s) == [3, 700, 9, oracle.tea(7, 300)]


# ---- B.2 packing, main.js:114-393
def test_packed_sizes(cornell):
    b = cornell.bindings()
    assert len(b[8]) == 1440 and len(b[3]) == 1024 and len(b[7]) == 80
    assert len(b[4]) == 64 and len(b[5]) == 5652 and len(b[6]) == 8428
    assert cornell.spectrum_index == {"white": 0, "green": 1, "red": 2, "light": 3, "dark": 4, "lightAlt": 5,
                                      "extinction": 6}


def test_packed_camera(cornell):
    assert cornell.camera.tolist() == [278, 273, -800, 0, 278, 273, 0, 0, 0, 1, 0, 1000, 1000,
                                       np.float32(0.7), 0, 0]


def test_packed_light(cornell):
    l = cornell.lights[0]
    assert l["category"] == 0
    assert l["data1"].tolist() == [213, 555, 227] and l["data2"].tolist() == [130, 0, 0]
    assert l["data3"].tolist() == [0, 0, 105] and l["data4"].tolist() == [3, 0, 1, 2]


def test_packed_primitives(cornell):
    p = cornell.primitives
    assert p["category"].tolist() == [0] * 16 + [1, 1]
    assert p["data4"][:, 3].tolist() == list(range(18))
    assert p[16]["data4"].tolist() == [4, 2, 0, 16]  # red diffuse sphere
    assert p[17]["data4"].tolist() == [4, 0, 2, 17]  # glass sphere
    assert p[17]["data1"].tolist() == [188, 240, 140] and p[17]["data2"].tolist() == [75, 75, 75]
    assert p[17]["data3"].tolist() == [0, 0, 0]
    # patch 6 has non-orthogonal edges (e1.e2 = 64): the u/v formula of CS:563-564 is kept as written
    assert float(np.dot(p[6]["data2"], p[6]["data3"])) == 64.0


def test_spectra_table(cornell):
    s = cornell.spectra
    assert hashlib.sha256(s.tobytes()).hexdigest().startswith("6ef5ac501722cf5c")
    u = s.view(np.uint32).reshape(7, 301)
    spots = {(0, 0): 0x3EAF9DB2, (0, 1): 0x3EBCAC08, (0, 150): 0x3F3CED91, (0, 300): 0x3F3CAC08,
             (1, 130): 0x3EF16873, (2, 250): 0x3F1C6A7F, (3, 0): 0x41700000, (3, 50): 0x41840000,
             (3, 150): 0x41866666, (3, 300): 0x41933333, (5, 0): 0x4065CBED, (6, 100): 0x3C23D70A,
             (6, 150): 0x3D6147AE}
    for (i, j), want in spots.items():
        assert int(u[i, j]) == want, (i, j)


def test_cie_table(cornell):
    c = cornell.cie
    assert hashlib.sha256(c.tobytes()).hexdigest().startswith("965e386c9f38c3f5")
    y = c.reshape(3, 471)[1]
    assert y[195] == 1.0
    # the constant 106.856895 of CS:422 is the sum of CIE_Y
    assert abs(float(y.astype(np.float64).sum()) - 106.856895) < 1e-4


# ---- B.3 primary rays, default scene, sample = 1
def test_camera_constants(oracle, cornell):
    sc = oracle.scene(cornell)
    eye, hor, ver, llc = sc.camera_constants()
    assert eye.tolist() == [278, 273, -800]
    np.testing.assert_allclose(hor, [-0.730057, 0, 0], rtol=2e-7, atol=0)
    np.testing.assert_allclose(ver, [0, 0.730057, 0], rtol=2e-7, atol=0)
    np.testing.assert_allclose(llc, [278.36502, 272.63498, -799], rtol=2e-7)


def test_primary_probes(oracle, cornell):
    sc = oracle.scene(cornell)
    p = sc.primary_probe(500, 500)
    assert abs(p["s"] - 0.50009197) < 1e-7 and abs(p["t"] - 0.500108421) < 1e-7
    np.testing.assert_allclose(p["dir"], [-6.10351562e-05, 9.15527344e-05, 1], rtol=1e-4, atol=1e-7)
    assert p["hit"] == 14 and abs(p["t_max"] - 1091.98901) < 2e-3
    p = sc.primary_probe(500, 50)
    assert p["hit"] == 1 and abs(p["t_max"] - 903.298401) < 2e-3
    p = sc.primary_probe(250, 500)
    assert p["hit"] == 3 and abs(p["t_max"] - 1377.37158) < 2e-3
    for px in ((0, 0), (999, 999)):
        p = sc.primary_probe(*px)
        assert p["hit"] == 0xFFFFFFFF and p["t_max"] == np.float32(2139095040.0)


HIST_1000 = {0: 90802, 1: 143764, 2: 5716, 3: 156040, 4: 146692, 5: 147195, 6: 4794, 7: 13, 8: 62747, 9: 6489,
             10: 3, 11: 12507, 14: 81349, 15: 24, 16: 6404, 17: 37961, 0xFFFFFFFF: 97500}
HIST_512 = {0: 23872, 1: 37936, 2: 1523, 3: 40866, 4: 38256, 5: 38663, 6: 1260, 7: 3, 8: 16321, 9: 1794, 10: 1,
            11: 3097, 14: 21469, 15: 6, 16: 1669, 17: 9946, 0xFFFFFFFF: 25462}


@pytest.mark.parametrize("size,want", [(512, HIST_512), (1000, HIST_1000)])
def test_primary_hit_histogram(oracle, size, want):
    """Survey pins were computed without fused multiply-adds; the numerics contract fuses dot/cross,
    so a few silhouette pixels may move between neighbours (SURVEY.md B.3 says as much)."""
    sc = oracle.scene(default_scene(size, size))
    idx, _ = sc.primary_hits(1)
    vals, counts = np.unique(idx, return_counts=True)
    got = {int(v): int(c) for v, c in zip(vals, counts)}
    assert set(got) == set(want)
    moved = sum(abs(got[k] - want[k]) for k in want)
    assert moved <= 12, (moved, {k: (got[k], want[k]) for k in want if got[k] != want[k]})
    assert sum(got.values()) == size * size


def test_detmath_is_accurate_against_double_libm(oracle):
    """include/crt_detmath.h is compiled into the oracle AND the kernels, so GPU == oracle says nothing about its accuracy
    (VERDICT r1, weak-1).  Against numpy's double-precision libm: sin and cos within 2 ulp on [0, 2 pi] (the only range
    the shader feeds them, CS:754-758: theta = 2 pi v, v in [0, 1)), exp within 2 ulp on [-87, 10] (CS:177, 438: arguments are
    -extinction x distance and -colour x exposure, never positive beyond the tone map's clamp; results below FLT_MIN flush
    towards zero), pow(x, 1/2.4) within 3e-7 relative on [0.0031308, 1] (the sRGB curve's domain, CS:449-465)."""
    rng = np.random.default_rng(42)

    def ulps(got, want64):
        want32 = want64.astype(np.float32)
        return np.abs(got.astype(np.float64) - want64) / np.spacing(np.abs(want32)).astype(np.float64)

    x = np.concatenate([rng.uniform(0.0, 2 * np.pi, 400_000), np.linspace(0, 2 * np.pi, 4097), [0.0, np.pi / 2, np.pi, 1.5 * np.pi]]).astype(np.float32)
    s, c, _, _ = oracle.detmath(x)
    x64 = x.astype(np.float64)
    # near the zeros of sin / cos an ulp of the RESULT is tiny while the argument itself carries half an ulp of 2 pi: bound
    # the absolute error there (2^-24) and the ulp error elsewhere
    for got, want in ((s, np.sin(x64)), (c, np.cos(x64))):
        big = np.abs(want) > 1e-2
        assert ulps(got[big], want[big]).max() <= 2.0
        assert np.abs(got[~big].astype(np.float64) - want[~big]).max() <= 2.0 ** -24
    xe = np.concatenate([rng.uniform(-87.0, 10.0, 400_000), np.linspace(-87, 10, 4097), [0.0, -0.0, 1.0, -1.0]]).astype(np.float32)
    _, _, e, _ = oracle.detmath(xe)
    assert ulps(e, np.exp(xe.astype(np.float64))).max() <= 2.0
    _, _, e, _ = oracle.detmath(np.array([-104.0, -200.0, -1e30, -np.inf], dtype=np.float32))
    assert np.all(e >= 0.0) and np.all(e < 1e-44)                       # far underflow: zero (or the smallest denormals), never garbage
    _, _, e, _ = oracle.detmath(np.array([np.nan], dtype=np.float32))
    assert np.isnan(e[0])
    xp = np.concatenate([rng.uniform(0.0031308, 1.0, 400_000), np.linspace(0.0031308, 1.0, 4097)]).astype(np.float32)
    _, _, _, p = oracle.detmath(xp)
    want = xp.astype(np.float64) ** (1.0 / 2.4)
    assert (np.abs(p.astype(np.float64) - want) / want).max() <= 3e-7
