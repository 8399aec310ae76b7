"""ctypes wrapper of the CPU oracle (oracle/crt_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product package never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "libcrt_oracle.so")
NUM_COUNTERS = 16


class Params(C.Structure):
    """struct crt_params of include/crt.h"""
    _fields_ = [("max_depth", C.c_uint32), ("grid_size", C.c_uint32), ("t_min", C.c_float), ("exposure", C.c_float),
                ("kernel", C.c_uint32), ("frames_per_launch", C.c_uint32), ("flags", C.c_uint32), ("wf_slots", C.c_uint32), ("reserved", C.c_uint32 * 4)]


def default_params(**kw) -> Params:
    p = Params(100, 16, 0.001, 2.2, 0, 0)
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def build() -> str:
    src = os.path.join(ORACLE_DIR, "crt_oracle.c")
    deps = [src, os.path.join(ROOT, "include", "crt.h"), os.path.join(ROOT, "include", "crt_detmath.h")]
    if not os.path.exists(LIB) or any(os.path.getmtime(d) > os.path.getmtime(LIB) for d in deps):
        subprocess.run(["make", "-C", ORACLE_DIR], check=True, capture_output=True)
    return LIB


class Oracle:
    def __init__(self, lib: C.CDLL):
        self.lib = lib
        vp, u32, f32p = C.c_void_p, C.c_uint32, C.POINTER(C.c_float)
        lib.orc_tea.restype = u32
        lib.orc_tea.argtypes = [u32, u32]
        lib.orc_pcg4d.argtypes = [C.POINTER(u32)]
        lib.orc_seed.argtypes = [u32, u32, u32, C.POINTER(u32)]
        lib.orc_scene_create.restype = vp
        lib.orc_scene_create.argtypes = [vp, u32, vp, u32, vp, vp, vp, u32, C.POINTER(Params), C.c_int]
        lib.orc_scene_destroy.argtypes = [vp]
        lib.orc_scene_camera_constants.argtypes = [vp, f32p]
        lib.orc_render.restype = C.c_int
        lib.orc_render.argtypes = [vp, u32, u32, vp, vp, vp, C.c_int, u32, u32, u32]
        lib.orc_resolve.argtypes = [vp, vp, u32, vp]
        lib.orc_trace_rays.argtypes = [vp, vp, vp, C.c_size_t, vp, vp, vp]
        lib.orc_primary_probe.argtypes = [vp, u32, u32, u32, f32p]
        lib.orc_primary_hits.argtypes = [vp, u32, vp, vp]
        lib.orc_detmath.argtypes = [vp, C.c_size_t, vp, vp, vp, vp]

    # ---- RNG
    def tea(self, a: int, b: int) -> int:
        return self.lib.orc_tea(a & 0xFFFFFFFF, b & 0xFFFFFFFF)

    def seed(self, px: int, py: int, sample: int):
        s = (C.c_uint32 * 4)()
        self.lib.orc_seed(px, py, sample, s)
        return s

    def rand_stream(self, px: int, py: int, sample: int, n: int):
        s = self.seed(px, py, sample)
        out = []
        for _ in range(n):
            self.lib.orc_pcg4d(s)
            out.append(s[0] & 0xFFFFFF)
        return out, list(s)

    def scene(self, ps, params: Params | None = None, use_bvh: bool = False) -> "OracleScene":
        return OracleScene(self, ps, params or default_params(), use_bvh)

    def detmath(self, x: np.ndarray):
        x = np.ascontiguousarray(x, dtype=np.float32)
        outs = [np.empty_like(x) for _ in range(4)]
        self.lib.orc_detmath(x.ctypes.data, x.size, *[o.ctypes.data for o in outs])
        return outs


class OracleScene:
    """Owns copies of the packed buffers for the lifetime of the C scene object."""

    def __init__(self, orc: Oracle, ps, params: Params, use_bvh: bool):
        self.orc = orc
        self.lib = orc.lib
        self.params = Params.from_buffer_copy(bytes(params))   # accepts the product's Params struct too (same layout)
        self.W, self.H = int(ps.camera[11]), int(ps.camera[12])
        self._prims = np.frombuffer(ps.bindings()[8], dtype=np.uint8).copy()
        self._lights = np.frombuffer(ps.bindings()[7], dtype=np.uint8).copy()
        self._cam = np.ascontiguousarray(ps.camera, dtype=np.float32).copy()
        self._cie = np.ascontiguousarray(ps.cie, dtype=np.float32).copy()
        self._spec = np.ascontiguousarray(ps.spectra, dtype=np.float32).copy()
        self.n_prims = self._prims.size // 80
        self.h = self.lib.orc_scene_create(self._prims.ctypes.data, self.n_prims, self._lights.ctypes.data,
                                           self._lights.size // 80, self._cam.ctypes.data, self._cie.ctypes.data,
                                           self._spec.ctypes.data, self._spec.size // 301, C.byref(self.params),
                                           1 if use_bvh else 0)

    def __del__(self):
        try:
            self.lib.orc_scene_destroy(self.h)
        except Exception:
            pass

    def camera_constants(self) -> np.ndarray:
        out = (C.c_float * 12)()
        self.lib.orc_scene_camera_constants(self.h, out)
        return np.array(out, dtype=np.float32).reshape(4, 3)  # eye, hor, ver, llc

    def render(self, n_frames: int, sample_start: int = 0, acc: np.ndarray | None = None, threads: int | None = None,
               rank: int = 0, world: int = 1, mode: int = 0, want_fb: bool = True):
        """Returns (acc[H,W,4] f32, fb[H,W,4] u8 or None, counters[16] u64)."""
        if acc is None:
            acc = np.zeros((self.H, self.W, 4), dtype=np.float32)
        else:
            acc = np.ascontiguousarray(acc, dtype=np.float32).copy()
        fb = np.zeros((self.H, self.W, 4), dtype=np.uint8) if want_fb else None
        cnt = np.zeros(NUM_COUNTERS, dtype=np.uint64)
        threads = threads or (os.cpu_count() or 1)
        self.lib.orc_render(self.h, sample_start, n_frames, acc.ctypes.data, fb.ctypes.data if want_fb else None,
                            cnt.ctypes.data, threads, rank, world, mode)
        return acc, fb, cnt

    def resolve(self, acc: np.ndarray, sample: int) -> np.ndarray:
        acc = np.ascontiguousarray(acc, dtype=np.float32)
        fb = np.zeros((self.H, self.W, 4), dtype=np.uint8)
        self.lib.orc_resolve(self.h, acc.ctypes.data, sample, fb.ctypes.data)
        return fb

    def trace_rays(self, rays: np.ndarray, exclude: np.ndarray | None = None):
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        n = rays.shape[0]
        idx = np.empty(n, dtype=np.uint32)
        t = np.empty(n, dtype=np.float32)
        cnt = np.zeros(NUM_COUNTERS, dtype=np.uint64)
        ex = None
        if exclude is not None:
            ex = np.ascontiguousarray(exclude, dtype=np.uint32)
        self.lib.orc_trace_rays(self.h, rays.ctypes.data, ex.ctypes.data if ex is not None else None, n,
                                idx.ctypes.data, t.ctypes.data, cnt.ctypes.data)
        return idx, t, cnt

    def primary_probe(self, px: int, py: int, sample: int = 1):
        out = (C.c_float * 8)()
        self.lib.orc_primary_probe(self.h, px, py, sample, out)
        a = np.array(out, dtype=np.float32)
        return dict(s=a[0], t=a[1], dir=a[2:5], hit=int(a[5:6].view(np.uint32)[0]), t_max=a[6])

    def primary_hits(self, sample: int = 1):
        idx = np.empty((self.H, self.W), dtype=np.uint32)
        t = np.empty((self.H, self.W), dtype=np.float32)
        self.lib.orc_primary_hits(self.h, sample, idx.ctypes.data, t.ctypes.data)
        return idx, t


def load() -> Oracle:
    return Oracle(C.CDLL(build()))
