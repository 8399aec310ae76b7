"""Scene packer: scene JSON -> the byte buffers the compute shader binds.

Host-side mirror of the packing code that lives inline in the reference's ``src/main.js``
(lines 114-393; ``src/sceneLoader.js`` itself is a 5-line stub that exports ``1``).  The output
is the reference's buffer contract, byte for byte:

=========  ==========================================  =====================
binding    contents                                     reference
=========  ==========================================  =====================
3          planar patches, 64-byte stride               main.js:139-209
4          camera, 16 x f32                             main.js:313-332
5          CIE X|Y|Z, 3 x 471 f32                        main.js:380-393
6          spectra resampled to 400..700 nm, n x 301    main.js:334-378
7          lights, 80-byte Primitive stride             main.js:255-296
8          primitives, 80-byte Primitive stride         main.js:114-137, 211-246
=========  ==========================================  =====================

Only numpy is used here; nothing in this module touches the GPU.
"""
from __future__ import annotations

import json
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np

LAMBDA_MIN = 400  # main.js:334
LAMBDA_MAX = 700  # main.js:335
N_LAMBDA = LAMBDA_MAX - LAMBDA_MIN + 1

#: main.js:166-170
TYPE_INDEX = {"diffuse": 0, "light": 1, "glass": 2}

#: category codes of the 80-byte Primitive record (ComputeShader.wgsl:525,584); 2 is this
#: project's triangle extension.
CATEGORY_PATCH, CATEGORY_SPHERE, CATEGORY_TRIANGLE = 0, 1, 2

PRIM_DTYPE = np.dtype(
    {
        "names": ["category", "data1", "data2", "data3", "data4"],
        "formats": ["<u4", ("<f4", 3), ("<f4", 3), ("<f4", 3), ("<u4", 4)],
        "offsets": [0, 16, 32, 48, 64],
        "itemsize": 80,
    }
)

PATCH_DTYPE = np.dtype(
    {
        "names": ["origin", "edge1", "edge2", "emission", "reflectance", "type", "index"],
        "formats": [("<f4", 3), ("<f4", 3), ("<f4", 3), "<u4", "<u4", "<u4", "<u4"],
        "offsets": [0, 16, 32, 44, 48, 52, 56],
        "itemsize": 64,
    }
)


def copy_records(a, dtype=None) -> np.ndarray:
    """A writable copy of a record array (or of its raw bytes) that keeps every byte, padding lanes included:
    ndarray.copy() / .tobytes() on a padded structured dtype go field by field and leave the padding undefined."""
    if isinstance(a, np.ndarray):
        dtype = dtype or a.dtype
        a = np.ascontiguousarray(a).view(np.uint8).tobytes()
    return np.frombuffer(bytearray(a), dtype=dtype)


@dataclass
class PackedScene:
    """The byte buffers of one scene, keyed like the shader's bindings."""

    width: int
    height: int
    planar_patches: np.ndarray  # binding 3
    camera: np.ndarray  # binding 4, float32[16]
    cie: np.ndarray  # binding 5, float32[3*471]
    spectra: np.ndarray  # binding 6, float32[n*301]
    lights: np.ndarray  # binding 7, PRIM_DTYPE[m]
    primitives: np.ndarray  # binding 8, PRIM_DTYPE[N]
    spectrum_index: Dict[str, int] = field(default_factory=dict)

    def bindings(self) -> Dict[int, bytes]:
        # the raw bytes of the record arrays: ndarray.tobytes() on a padded structured dtype copies field by field and
        # leaves the padding lanes of the copy uninitialised, which made the uploaded bytes (and any digest of them)
        # differ from run to run
        def raw(a: np.ndarray) -> bytes:
            return np.ascontiguousarray(a).view(np.uint8).tobytes()
        return {
            3: raw(self.planar_patches),
            4: self.camera.tobytes(),
            5: self.cie.tobytes(),
            6: self.spectra.tobytes(),
            7: raw(self.lights),
            8: raw(self.primitives),
        }

    @property
    def n_spectra(self) -> int:
        return self.spectra.size // N_LAMBDA


def _lerp(a: float, b: float, t: float) -> float:
    """main.js:626 -- evaluated in double like JavaScript numbers."""
    return a * (1 - t) + b * t


def sample_spectrum(spectrum: dict, lam: float) -> float:
    """main.js:340-356 ``sampleSpectrum``: piecewise-linear lookup, clamped at both ends."""
    wl = spectrum["wavelength"]
    val = spectrum["value"]
    index = next((i for i, e in enumerate(wl) if e >= lam), -1)  # Array.findIndex
    start_index = max(index - 1, 0)
    end_index = min(index, len(wl) - 1)
    # JS: a negative index reads `undefined`; findIndex returns -1 only past the last
    # wavelength, where Math.min(-1, n-1) = -1 -> undefined -> NaN.  The shipped scenes never
    # reach that (every table ends at >= 700 nm); clamp so the packer stays total.
    if index < 0:
        start_index = end_index = len(wl) - 1
    start, end = val[start_index], val[end_index]
    start_lambda, end_lambda = wl[start_index], wl[end_index]
    if start_lambda == end_lambda:
        return start
    return _lerp(start, end, (lam - start_lambda) / (end_lambda - start_lambda))


def pack_spectra(spectra: dict) -> np.ndarray:
    """main.js:358-378: every spectrum (JS object key order) at 400..700 nm in 1-nm steps."""
    out = np.empty(len(spectra) * N_LAMBDA, dtype=np.float64)
    k = 0
    for name in spectra:  # json.load keeps the file's key order, like Object.keys
        for i in range(N_LAMBDA):
            out[k] = sample_spectrum(spectra[name], LAMBDA_MIN + i)
            k += 1
    return out.astype("<f4")


def pack_cie(cie: dict) -> np.ndarray:
    """main.js:380-393: CIE_X ++ CIE_Y ++ CIE_Z as f32."""
    return np.concatenate([np.asarray(cie[k], dtype=np.float64) for k in ("CIE_X", "CIE_Y", "CIE_Z")]).astype("<f4")


def pack_camera(cam: dict) -> np.ndarray:
    """main.js:313-324: [eye,0, lookat,0, up,width, height,focalLength,0,0]."""
    return np.asarray(
        [*cam["eye"], 0, *cam["lookat"], 0, *cam["up"], cam["width"], cam["height"], cam["focalLength"], 0, 0],
        dtype="<f4",
    )


def load_scene(scene: dict, cie: dict, extra_triangles: Optional[np.ndarray] = None,
               triangle_reflectance: str = "white") -> PackedScene:
    """Pack a scene dict (the schema of ``src/scenes/cornell.json``) and the CIE tables.

    ``extra_triangles`` (float32[n,3,3], vertices v0,v1,v2) appends category-2 primitives after
    the patches and spheres; that is this project's extension for the large-scene configs and
    has no counterpart in the reference's packer.
    """
    key_index = {name: i for i, name in enumerate(scene["spectra"])}  # main.js:161-164
    patches = scene["objects"].get("patches", [])
    spheres = scene["objects"].get("spheres", [])

    # main.js:116-137: primitives = patches (index i) ++ spheres (index n_patches + i)
    n_tri = 0 if extra_triangles is None else int(extra_triangles.shape[0])
    prims = np.zeros(len(patches) + len(spheres) + n_tri, dtype=PRIM_DTYPE)
    pp = np.zeros(len(patches), dtype=PATCH_DTYPE)
    light_rows: List[int] = []
    for i, p in enumerate(patches):
        # main.js:172-209 (binding 3) and main.js:224-233 (binding 8)
        pp[i]["origin"], pp[i]["edge1"], pp[i]["edge2"] = p["origin"], p["edge1"], p["edge2"]
        pp[i]["emission"] = key_index[p["emission"]]
        pp[i]["reflectance"] = key_index[p["reflectance"]]
        pp[i]["type"] = TYPE_INDEX[p["type"]]
        pp[i]["index"] = i
        r = prims[i]
        r["category"] = CATEGORY_PATCH
        r["data1"], r["data2"], r["data3"] = p["origin"], p["edge1"], p["edge2"]
        r["data4"] = (key_index[p["emission"]], key_index[p["reflectance"]], TYPE_INDEX[p["type"]], i)
        if p["type"] == "light":
            light_rows.append(i)
    base = len(patches)
    for i, s in enumerate(spheres):
        # main.js:234-244: centre in data1, (r, r, r) in data2, data3 left zero
        r = prims[base + i]
        r["category"] = CATEGORY_SPHERE
        r["data1"] = s["center"]
        r["data2"] = (s["radius"],) * 3
        r["data4"] = (key_index[s["emission"]], key_index[s["reflectance"]], TYPE_INDEX[s["type"]], base + i)
        if s["type"] == "light":
            light_rows.append(base + i)
    base += len(spheres)
    if n_tri:
        tri = np.asarray(extra_triangles, dtype="<f4")
        sl = slice(base, base + n_tri)
        prims["category"][sl] = CATEGORY_TRIANGLE
        prims["data1"][sl] = tri[:, 0]
        prims["data2"][sl] = tri[:, 1] - tri[:, 0]
        prims["data3"][sl] = tri[:, 2] - tri[:, 0]
        prims["data4"][sl, 0] = key_index["dark"] if "dark" in key_index else 0
        prims["data4"][sl, 1] = key_index[triangle_reflectance]
        prims["data4"][sl, 2] = TYPE_INDEX["diffuse"]
        prims["data4"][sl, 3] = np.arange(base, base + n_tri, dtype=np.uint32)

    # main.js:255-296: lights = primitives whose type is "light", written with the patch fields
    # (origin/edge1/edge2); a sphere light has none of those in the reference (they would be
    # `undefined`), so only patch lights are meaningful.
    # (copied as raw 80-byte rows so the padding lanes stay zero and the bytes are deterministic)
    raw = prims.view(np.uint8).reshape(-1, PRIM_DTYPE.itemsize)
    lights = raw[light_rows].copy().reshape(-1).view(PRIM_DTYPE)

    cam = scene["camera"]
    return PackedScene(
        width=int(cam["width"]),
        height=int(cam["height"]),
        planar_patches=pp,
        camera=pack_camera(cam),
        cie=pack_cie(cie),
        spectra=pack_spectra(scene["spectra"]),
        lights=lights,
        primitives=prims,
        spectrum_index=key_index,
    )


def load_scene_files(scene_path: str, cie_path: str, **kw) -> PackedScene:
    with open(scene_path) as f:
        scene = json.load(f)
    with open(cie_path) as f:
        cie = json.load(f)
    return load_scene(scene, cie, **kw)
