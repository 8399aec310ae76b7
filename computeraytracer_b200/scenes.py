"""Scenes of the BASELINE.json configs, as packed binding buffers.

The default scene is the reference's ``src/scenes/cornell.json`` run through the packer
(``scene_loader.py``) and stored in ``data/default_scene.npz`` by
``tests/golden/make_fixtures.py`` (the GPU box has no copy of the reference).  The large
scenes are procedural because neither the reference nor this image ships a mesh: they reuse
the Cornell walls, light, spectra and camera (so ``t_min = 0.001`` keeps its meaning at the
555-unit scale) and add category-2 triangles.
"""
from __future__ import annotations

import os
from typing import Optional

import numpy as np

from .scene_loader import (CATEGORY_TRIANGLE, PATCH_DTYPE, PRIM_DTYPE, PackedScene, copy_records)

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "default_scene.npz")


def default_scene(width: Optional[int] = None, height: Optional[int] = None) -> PackedScene:
    """cornell.json (16 patches, 2 spheres, 1 light); 1000x1000 unless overridden."""
    z = np.load(_DATA)
    names = [str(s) for s in z["spectrum_names"]]
    cam = z["camera"].astype("<f4").copy()
    if width is not None:
        cam[11] = width
    if height is not None:
        cam[12] = height
    return PackedScene(
        width=int(cam[11]),
        height=int(cam[12]),
        planar_patches=copy_records(z["planar_patches"].tobytes(), PATCH_DTYPE),
        camera=cam,
        cie=z["cie"].astype("<f4"),
        spectra=z["spectra"].astype("<f4"),
        lights=copy_records(z["lights"].tobytes(), PRIM_DTYPE),
        primitives=copy_records(z["primitives"].tobytes(), PRIM_DTYPE),
        spectrum_index={n: i for i, n in enumerate(names)},
    )


def _with_triangles(base: PackedScene, keep: int, tris: np.ndarray, reflectance: str = "white") -> PackedScene:
    """First ``keep`` primitives of ``base`` (walls + light) followed by triangles."""
    n = tris.shape[0]
    prims = np.zeros(keep + n, dtype=PRIM_DTYPE)
    prims[:keep] = base.primitives[:keep]
    sl = slice(keep, keep + n)
    t = np.asarray(tris, dtype="<f4")
    prims["category"][sl] = CATEGORY_TRIANGLE
    prims["data1"][sl] = t[:, 0]
    prims["data2"][sl] = t[:, 1] - t[:, 0]
    prims["data3"][sl] = t[:, 2] - t[:, 0]
    prims["data4"][sl, 0] = base.spectrum_index["dark"]
    prims["data4"][sl, 1] = base.spectrum_index[reflectance]
    prims["data4"][sl, 2] = 0  # diffuse
    prims["data4"][sl, 3] = np.arange(keep, keep + n, dtype=np.uint32)
    return PackedScene(
        width=base.width, height=base.height,
        planar_patches=copy_records(base.planar_patches[:keep]),
        camera=base.camera.copy(), cie=base.cie, spectra=base.spectra,
        lights=copy_records(base.lights), primitives=prims, spectrum_index=dict(base.spectrum_index),
    )


def _value_noise3(p: np.ndarray, seed: int, octaves: int = 4, base_freq: float = 3.0) -> np.ndarray:
    """Seeded 3-D value noise in [-1, 1] (smoothstep-interpolated random lattice, summed octaves)."""
    rng = np.random.default_rng(seed)
    L = 64
    lattice = rng.random((L, L, L), dtype=np.float64) * 2.0 - 1.0
    out = np.zeros(p.shape[0], dtype=np.float64)
    amp, freq, norm = 1.0, base_freq, 0.0
    for _ in range(octaves):
        q = p * freq + 17.0
        i0 = np.floor(q).astype(np.int64)
        f = q - i0
        f = f * f * (3.0 - 2.0 * f)
        acc = np.zeros(p.shape[0], dtype=np.float64)
        for dx in (0, 1):
            wx = f[:, 0] if dx else 1.0 - f[:, 0]
            for dy in (0, 1):
                wy = f[:, 1] if dy else 1.0 - f[:, 1]
                for dz in (0, 1):
                    wz = f[:, 2] if dz else 1.0 - f[:, 2]
                    acc += wx * wy * wz * lattice[(i0[:, 0] + dx) % L, (i0[:, 1] + dy) % L, (i0[:, 2] + dz) % L]
        out += amp * acc
        norm += amp
        amp *= 0.5
        freq *= 2.0
    return out / norm


def displaced_sphere_triangles(n_lon: int = 1024, n_lat: int = 512, seed: int = 1,
                               center=(278.0, 200.0, 280.0), radius: float = 150.0,
                               displacement: float = 0.25) -> np.ndarray:
    """Closed displaced sphere, 2*n_lon*n_lat triangles (1,048,576 at the defaults).

    A longitude/latitude grid; the two triangle rows touching the poles contain zero-area
    triangles, which never intersect anything (det == 0) and are kept so the count is exact.
    """
    lon = np.linspace(0.0, 2.0 * np.pi, n_lon, endpoint=False)
    lat = np.linspace(0.0, np.pi, n_lat + 1)
    LON, LAT = np.meshgrid(lon, lat, indexing="xy")  # (n_lat+1, n_lon)
    d = np.stack([np.sin(LAT) * np.cos(LON), np.cos(LAT), np.sin(LAT) * np.sin(LON)], axis=-1).reshape(-1, 3)
    r = radius * (1.0 + displacement * _value_noise3(d, seed))
    verts = (np.asarray(center, dtype=np.float64)[None, :] + d * r[:, None]).astype("<f4")
    verts = verts.reshape(n_lat + 1, n_lon, 3)
    a = verts[:-1, :, :]
    b = np.roll(verts, -1, axis=1)[:-1, :, :]
    c = verts[1:, :, :]
    e = np.roll(verts, -1, axis=1)[1:, :, :]
    t1 = np.stack([a, c, b], axis=2)  # (n_lat, n_lon, 3 verts, 3)
    t2 = np.stack([b, c, e], axis=2)
    return np.concatenate([t1.reshape(-1, 3, 3), t2.reshape(-1, 3, 3)], axis=0)


def mesh_scene(width: int = 1920, height: int = 1080, n_lon: int = 1024, n_lat: int = 512, seed: int = 1) -> PackedScene:
    """BASELINE config 3: Cornell walls + light (patches 0..5) and a displaced-sphere mesh."""
    base = default_scene(width, height)
    tris = displaced_sphere_triangles(n_lon, n_lat, seed)
    return _with_triangles(base, 6, tris)


def soup_triangles(n: int, seed: int = 5, lo: float = 5.0, hi: float = 550.0, edge: float = 6.0) -> np.ndarray:
    """n random triangles: centroid uniform in the box interior, edge vectors uniform in a cube of side ``edge``."""
    rng = np.random.default_rng(seed)  # PCG64
    c = rng.uniform(lo, hi, size=(n, 3))
    e1 = rng.uniform(-edge / 2, edge / 2, size=(n, 3))
    e2 = rng.uniform(-edge / 2, edge / 2, size=(n, 3))
    v0 = c - (e1 + e2) / 3.0
    return np.stack([v0, v0 + e1, v0 + e2], axis=1).astype("<f4")


def soup_scene(width: int = 1920, height: int = 1080, n: int = 10_000_000, seed: int = 5) -> PackedScene:
    """BASELINE config 5: Cornell walls + light and an n-triangle random soup."""
    base = default_scene(width, height)
    return _with_triangles(base, 6, soup_triangles(n, seed))


def with_triangle_lights(ps: PackedScene, tris: np.ndarray, emission: str = "light") -> PackedScene:
    """``ps`` plus emitting triangles: each is appended to binding 8 (category 2, material light) and to binding 7 as a light
    record (origin, edge1, edge2 as the reference's patch lights, MJ:255-296; data4 = (emission, 0, light, primitive index)).
    An extension: the reference has one patch light.  Its shader samples a light's parallelogram origin + u e1 + v e2
    (CS:341-355), so half of the samples towards a triangle light land beside it and contribute nothing -- the semantics as
    written, which the oracle and the kernels share."""
    t = np.asarray(tris, dtype="<f4")
    k, n0, l0 = t.shape[0], ps.primitives.size, ps.lights.size
    prims = np.zeros(n0 + k, dtype=PRIM_DTYPE)
    prims.view(np.uint8)[: n0 * 80] = copy_records(ps.primitives).view(np.uint8)
    lights = np.zeros(l0 + k, dtype=PRIM_DTYPE)
    lights.view(np.uint8)[: l0 * 80] = copy_records(ps.lights).view(np.uint8)
    e = ps.spectrum_index[emission]
    for rec, sl in ((prims, slice(n0, n0 + k)), (lights, slice(l0, l0 + k))):
        rec["category"][sl] = CATEGORY_TRIANGLE
        rec["data1"][sl] = t[:, 0]
        rec["data2"][sl] = t[:, 1] - t[:, 0]
        rec["data3"][sl] = t[:, 2] - t[:, 0]
        rec["data4"][sl, 0] = e
        rec["data4"][sl, 1] = 0
        rec["data4"][sl, 2] = 1  # light
        rec["data4"][sl, 3] = np.arange(n0, n0 + k, dtype=np.uint32)
    return PackedScene(width=ps.width, height=ps.height, planar_patches=copy_records(ps.planar_patches), camera=ps.camera.copy(), cie=ps.cie,
                       spectra=ps.spectra, lights=lights, primitives=prims, spectrum_index=dict(ps.spectrum_index))


def camera_rays(ps: PackedScene, n: int, seed: int = 0) -> np.ndarray:
    """n random rays (n x 8 f32: origin, -, direction, -) through the scene volume, for hit-parity tests.

    Half start at the camera eye towards random points of the box, half start inside the box
    with uniformly random directions (the incoherent-bounce case).
    """
    rng = np.random.default_rng(seed)
    rays = np.zeros((n, 8), dtype="<f4")
    h = n // 2
    eye = ps.camera[0:3].astype(np.float64)
    tgt = rng.uniform(0.0, 555.0, size=(h, 3))
    d = tgt - eye[None, :]
    rays[:h, 0:3] = eye
    rays[:h, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    o = rng.uniform(1.0, 554.0, size=(n - h, 3))
    v = rng.normal(size=(n - h, 3))
    rays[h:, 0:3] = o
    rays[h:, 4:7] = v / np.linalg.norm(v, axis=1, keepdims=True)
    return rays
