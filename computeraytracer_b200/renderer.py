"""Host-side mirror of the reference's ``src/main.js`` over the libcrt C ABI.

``Main()`` in the reference (main.js:7-624) creates the device, uploads the packed scene into
nine bindings, and then runs ``frame()`` forever: ``sample++`` followed by one compute dispatch
(main.js:597-611).  :class:`Renderer` is the same sequence against ``libcrt.so``:

* ``Renderer(device)``            -- requestAdapter/requestDevice + pipeline (main.js:8-9, 535-543)
* ``set_scene(packed)``           -- createBuffer/unmap of bindings 3..8 (main.js:139-393)
* ``frame(n)``                    -- n x frame() (main.js:584-621) without the blit
* ``accumulator()/framebuffer()`` -- readback of bindings 1 / 0 (new: the reference never reads back)

The reference's host language is JavaScript; ``node`` is not available in this image, so this
module (Python + ctypes, no torch) is the tested host layer and ``js/`` holds the N-API source a
Node user would build.  There is no CPU fallback: importing works anywhere,
but creating a Renderer without ``libcrt.so`` and a B200 raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

from .scene_loader import PackedScene

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CRT_LIBCRT") or os.path.join(_HERE, "libcrt.so")   # CRT_LIBCRT: a tuning build (build.build_variant)

NUM_COUNTERS = 16
COUNTER_NAMES = ["paths", "rays_ext", "rays_shadow", "prim_tests", "node_visits", "tri_tests", "rand",
                 "term_miss", "term_rr", "term_light", "term_depth", "depth_sum", "launches", "wf_iterations", "bvh_rays",
                 "max_ray_nodes"]
SHARD_NONE, SHARD_TILES, SHARD_SPLIT_SPP = 0, 1, 2
KERNEL_AUTO, KERNEL_MEGA, KERNEL_WAVEFRONT = 0, 1, 2
FLAG_STATS, FLAG_KERNEL_TIMES, FLAG_WF_SPLIT = 1, 2, 4
KERNEL_CLASSES = ["wf_logic", "wf_traverse", "wf_accumulate", "pt_megakernel", "wf_regen", "wf_shade", "wf_partition", "wf_tail"]

# crt_params.wf_tuning / bvh_tuning bits (include/crt.h)
WF_LIVE_ONE_LIST, WF_LIVE_ROUTED, WF_LIVE_CLASS = 1, 2, 3
WF_DEFER_OFF, WF_DEFER_ON, WF_HOST_LOOP, WF_ONE_LANE, WF_TWO_LANES_ALWAYS = 4, 8, 16, 32, 64
WF_TAIL_SHIFT, WF_TAIL_OFF = 8, 31 << 8   # bits 8-12: v = 1..30 -> wf_tail takes over at <= 2^(v-1) live slots; 31: never
BVH_EXACT_SLOTS, BVH_BUILDER_HOST_SAH, BVH_BUILDER_DEVICE = 0x200, 0x400, 0x800
COMM_ID_BYTES = 128
NOALLOC_MAX_RAYS = 65536   # CRT_NOALLOC_MAX_RAYS

STATUS = {0: "CRT_OK", 1: "CRT_ERR_ARG", 2: "CRT_ERR_BINDING", 3: "CRT_ERR_SIZE", 4: "CRT_ERR_STATE",
          5: "CRT_ERR_CUDA", 6: "CRT_ERR_OOM", 7: "CRT_ERR_NODEVICE", 8: "CRT_ERR_NCCL"}


class CrtError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"{STATUS.get(code, code)}: {msg}")
        self.code = code


class Params(C.Structure):
    """``struct crt_params`` (include/crt.h)."""
    _fields_ = [("max_depth", C.c_uint32), ("grid_size", C.c_uint32), ("t_min", C.c_float), ("exposure", C.c_float),
                ("kernel", C.c_uint32), ("frames_per_launch", C.c_uint32), ("flags", C.c_uint32), ("wf_slots", C.c_uint32),
                ("wf_tuning", C.c_uint32), ("bvh_tuning", C.c_uint32), ("bvh_prim_cost", C.c_float), ("slow_ray_nodes", C.c_uint32)]


_lib = None


def load_library(path: Optional[str] = None) -> C.CDLL:
    """dlopen libcrt.so and declare the prototypes of include/crt.h.  Raises if it is missing."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or LIB_PATH
    if not os.path.exists(path):
        raise FileNotFoundError(f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                                "(there is no CPU fallback)")
    lib = C.CDLL(path)
    vp, u32, sz = C.c_void_p, C.c_uint32, C.c_size_t
    lib.crt_version.restype = C.c_char_p
    lib.crt_last_error.restype = C.c_char_p
    lib.crt_last_error.argtypes = [vp]
    lib.crt_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.crt_destroy.argtypes = [vp]
    lib.crt_destroy.restype = None
    lib.crt_set_binding.argtypes = [vp, u32, vp, sz]
    lib.crt_reset.argtypes = [vp]
    lib.crt_default_params.argtypes = [C.POINTER(Params)]
    lib.crt_default_params.restype = None
    lib.crt_set_params.argtypes = [vp, C.POINTER(Params)]
    lib.crt_get_params.argtypes = [vp, C.POINTER(Params)]
    lib.crt_set_shard.argtypes = [vp, u32, u32, u32]
    lib.crt_dispatch_frames.argtypes = [vp, u32]
    lib.crt_synchronize.argtypes = [vp]
    lib.crt_last_dispatch_ms.argtypes = [vp, C.POINTER(C.c_float)]
    lib.crt_read_binding.argtypes = [vp, u32, vp, sz]
    lib.crt_resolve_framebuffer.argtypes = [vp]
    lib.crt_device_ptr.argtypes = [vp, u32, C.POINTER(vp), C.POINTER(sz)]
    lib.crt_stream.argtypes = [vp, C.POINTER(vp)]
    lib.crt_set_stream.argtypes = [vp, vp]
    lib.crt_trace_rays.argtypes = [vp, vp, vp, sz, vp, vp]
    lib.crt_trace_rays_noalloc.argtypes = [vp, vp, vp, sz, vp, vp]
    lib.crt_get_counters.argtypes = [vp, vp, sz]
    lib.crt_reset_counters.argtypes = [vp]
    lib.crt_get_kernel_times.argtypes = [vp, vp, vp, sz]
    lib.crt_scene_stats.argtypes = [vp, vp, sz]
    lib.crt_test_detmath.argtypes = [vp, vp, sz, vp, vp, vp, vp]
    lib.crt_measure_fp32_tflops.argtypes = [vp, C.POINTER(C.c_double)]
    lib.crt_create_multi.argtypes = [C.POINTER(C.c_int), C.c_int, u32, C.POINTER(vp)]
    lib.crt_comm_unique_id.argtypes = [vp, sz]
    lib.crt_comm_init.argtypes = [vp, vp, sz, u32, u32, u32]
    lib.crt_comm_destroy.argtypes = [vp]
    lib.crt_combine.argtypes = [vp, u32]
    lib.crt_last_combine.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_uint64)]
    lib.crt_read_bvh.argtypes = [vp, vp, sz, vp, sz, vp, vp]
    lib.crt_host_build_bvh.argtypes = [vp, sz, C.c_int, vp, sz, vp, sz, vp, vp, u32, C.c_float]
    if path == LIB_PATH:
        _lib = lib
    return lib


def default_params(**kw) -> Params:
    p = Params()
    load_library().crt_default_params(C.byref(p))
    for k, v in kw.items():
        setattr(p, k, v)
    return p


class Renderer:
    """One libcrt context: one GPU (``Renderer(device)``) or, with ``devices=[...]``, a single-process group of GPUs
    (``crt_create_multi``) that shards tiles or frames and combines them on its first member at readback."""

    def __init__(self, device: int = 0, lib: Optional[C.CDLL] = None, devices=None, shard_mode: int = SHARD_SPLIT_SPP):
        self.lib = lib or load_library()
        h = C.c_void_p()
        if devices is not None:
            arr = (C.c_int * len(devices))(*devices)
            rc = self.lib.crt_create_multi(arr, len(devices), shard_mode, C.byref(h))
            device = devices[0] if len(devices) else 0
        else:
            rc = self.lib.crt_create(device, C.byref(h))
        if rc:
            raise CrtError(rc, f"crt_create(device={devices if devices is not None else device}) failed -- libcrt needs a B200 (sm_100); there is no CPU fallback")
        self.h = h
        self.device = device
        self.devices = list(devices) if devices is not None else [device]
        self.width = self.height = 0

    # ---- lifetime
    def close(self) -> None:
        if getattr(self, "h", None):
            self.lib.crt_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc: int) -> None:
        if rc:
            raise CrtError(rc, self.lib.crt_last_error(self.h).decode())

    # ---- inputs
    def set_binding(self, binding: int, data) -> None:
        buf = data if isinstance(data, (bytes, bytearray)) else np.ascontiguousarray(data).tobytes()
        self._ck(self.lib.crt_set_binding(self.h, binding, buf, len(buf)))

    def set_scene(self, ps: PackedScene) -> None:
        """Upload bindings 3..8 in the reference's layouts (main.js:139-393)."""
        for b, data in ps.bindings().items():
            self.set_binding(b, data)
        self.width, self.height = int(ps.camera[11]), int(ps.camera[12])

    def set_camera(self, camera16: np.ndarray) -> None:
        cam = np.ascontiguousarray(camera16, dtype="<f4")
        self.set_binding(4, cam)
        self.width, self.height = int(cam[11]), int(cam[12])

    def set_params(self, params: Optional[Params] = None, **kw) -> Params:
        p = params or self.get_params()
        for k, v in kw.items():
            setattr(p, k, v)
        self._ck(self.lib.crt_set_params(self.h, C.byref(p)))
        return p

    def get_params(self) -> Params:
        p = Params()
        self._ck(self.lib.crt_get_params(self.h, C.byref(p)))
        return p

    def enable_stats(self, on: bool = True) -> None:
        """Detailed device counters (primitive tests, node visits, rand calls, termination mix)."""
        p = self.get_params()
        p.flags = (p.flags | FLAG_STATS) if on else (p.flags & ~FLAG_STATS)
        self._ck(self.lib.crt_set_params(self.h, C.byref(p)))

    def enable_kernel_times(self, on: bool = True) -> None:
        """CUDA events between the hot-path launches, read with :meth:`kernel_times`."""
        p = self.get_params()
        p.flags = (p.flags | FLAG_KERNEL_TIMES) if on else (p.flags & ~FLAG_KERNEL_TIMES)
        self._ck(self.lib.crt_set_params(self.h, C.byref(p)))

    def set_shard(self, rank: int, world: int, mode: int) -> None:
        self._ck(self.lib.crt_set_shard(self.h, rank, world, mode))

    # ---- multi-GPU, one process per GPU: NCCL behind the C ABI (the host layer only carries the 128-byte id)
    def comm_unique_id(self) -> bytes:
        buf = C.create_string_buffer(COMM_ID_BYTES)
        rc = self.lib.crt_comm_unique_id(buf, COMM_ID_BYTES)
        if rc:
            raise CrtError(rc, "crt_comm_unique_id failed (is libnccl.so.2 loadable?)")
        return buf.raw

    def comm_init(self, comm_id: bytes, rank: int, world: int, mode: int) -> None:
        self._ck(self.lib.crt_comm_init(self.h, comm_id, len(comm_id), rank, world, mode))

    def comm_destroy(self) -> None:
        self._ck(self.lib.crt_comm_destroy(self.h))

    def combine(self, dst: int = 0) -> None:
        """Collective: split-spp partial sums are reduced, owned tiles gathered, to rank ``dst`` (a group: its member 0),
        which also rebuilds binding 0.  Repeatable; asynchronous on the context's stream."""
        self._ck(self.lib.crt_combine(self.h, dst))

    def last_combine(self):
        """(device ms of the most recent combine, bytes this rank sent)"""
        ms, nb = C.c_float(), C.c_uint64()
        self._ck(self.lib.crt_last_combine(self.h, C.byref(ms), C.byref(nb)))
        return float(ms.value), int(nb.value)

    def reset(self) -> None:
        self._ck(self.lib.crt_reset(self.h))

    # ---- the frame loop
    def frame(self, n: int = 1) -> None:
        """n x frame() of main.js:584-621 (sample++ then one sample per pixel); asynchronous."""
        self._ck(self.lib.crt_dispatch_frames(self.h, n))

    def synchronize(self) -> None:
        self._ck(self.lib.crt_synchronize(self.h))

    def last_dispatch_ms(self) -> float:
        ms = C.c_float()
        self._ck(self.lib.crt_last_dispatch_ms(self.h, C.byref(ms)))
        return float(ms.value)

    # ---- outputs
    @property
    def sample(self) -> int:
        v = C.c_uint32()
        self._ck(self.lib.crt_read_binding(self.h, 2, C.byref(v), 4))
        return int(v.value)

    def accumulator(self, out: Optional[np.ndarray] = None) -> np.ndarray:
        """Binding 1: float32[H, W, 4] XYZ sums (lane 3 is padding)."""
        if out is None:
            out = np.empty((self.height, self.width, 4), dtype=np.float32)
        self._ck(self.lib.crt_read_binding(self.h, 1, out.ctypes.data, out.nbytes))
        return out

    def framebuffer(self, out: Optional[np.ndarray] = None) -> np.ndarray:
        """Binding 0: uint8[H, W, 4] RGBA, top-left origin."""
        if out is None:
            out = np.empty((self.height, self.width, 4), dtype=np.uint8)
        self._ck(self.lib.crt_read_binding(self.h, 0, out.ctypes.data, out.nbytes))
        return out

    def resolve_framebuffer(self) -> None:
        self._ck(self.lib.crt_resolve_framebuffer(self.h))

    def device_ptr(self, binding: int):
        p, n = C.c_void_p(), C.c_size_t()
        self._ck(self.lib.crt_device_ptr(self.h, binding, C.byref(p), C.byref(n)))
        return int(p.value), int(n.value)

    def stream(self) -> int:
        s = C.c_void_p()
        self._ck(self.lib.crt_stream(self.h, C.byref(s)))
        return int(s.value or 0)

    def set_stream(self, cuda_stream: int) -> None:
        """Launch on a caller-owned CUDA stream (0/None: the context's own)."""
        self._ck(self.lib.crt_set_stream(self.h, C.c_void_p(cuda_stream or None)))

    # ---- hooks
    def trace_rays(self, rays: np.ndarray, exclude: Optional[np.ndarray] = None):
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        n = rays.shape[0]
        idx = np.empty(n, dtype=np.uint32)
        t = np.empty(n, dtype=np.float32)
        ex = None if exclude is None else np.ascontiguousarray(exclude, dtype=np.uint32)
        self._ck(self.lib.crt_trace_rays(self.h, rays.ctypes.data, ex.ctypes.data if ex is not None else None, n,
                                         idx.ctypes.data, t.ctypes.data))
        return idx, t

    def trace_rays_noalloc(self, rays: np.ndarray, exclude: Optional[np.ndarray] = None):
        """``crt_trace_rays_noalloc``: as :meth:`trace_rays`, but the call allocates nothing on the device (at most
        ``NOALLOC_MAX_RAYS`` rays; the scene must be resident)."""
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        n = rays.shape[0]
        idx = np.empty(n, dtype=np.uint32)
        t = np.empty(n, dtype=np.float32)
        ex = None if exclude is None else np.ascontiguousarray(exclude, dtype=np.uint32)
        self._ck(self.lib.crt_trace_rays_noalloc(self.h, rays.ctypes.data, ex.ctypes.data if ex is not None else None, n,
                                                 idx.ctypes.data, t.ctypes.data))
        return idx, t

    def counters(self) -> dict:
        out = np.zeros(NUM_COUNTERS, dtype=np.uint64)
        self._ck(self.lib.crt_get_counters(self.h, out.ctypes.data, NUM_COUNTERS))
        return {k: int(out[i]) for i, k in enumerate(COUNTER_NAMES)}

    def kernel_times(self) -> dict:
        """{kernel class: (summed device ms, launches)} since the last reset_counters()."""
        ms = np.zeros(len(KERNEL_CLASSES), dtype=np.float64)
        n = np.zeros(len(KERNEL_CLASSES), dtype=np.uint64)
        self._ck(self.lib.crt_get_kernel_times(self.h, ms.ctypes.data, n.ctypes.data, len(KERNEL_CLASSES)))
        return {k: (float(ms[i]), int(n[i])) for i, k in enumerate(KERNEL_CLASSES)}

    def reset_counters(self) -> None:
        self._ck(self.lib.crt_reset_counters(self.h))

    def scene_stats(self) -> dict:
        out = np.zeros(10, dtype=np.uint64)
        self._ck(self.lib.crt_scene_stats(self.h, out.ctypes.data, 10))
        keys = ["n_prims", "n_linear", "n_triangles", "bvh_nodes", "bvh_bytes", "tri_bytes", "build_us", "bvh_depth", "tree_us"]
        d = {k: int(out[i]) for i, k in enumerate(keys)}
        d["builder"] = {0: "none", 1: "host_sah", 2: "device_ploc"}[int(out[9]) & 0xFF]
        d["shadow_closest"] = bool(int(out[9]) & 0x100)
        return d

    def measure_fp32_tflops(self) -> float:
        """Measured FP32 FMA rate of the device (TFLOP/s): the yardstick of the linear-list (FP32-bound) regime."""
        v = C.c_double()
        self._ck(self.lib.crt_measure_fp32_tflops(self.h, C.byref(v)))
        return float(v.value)

    def read_bvh(self):
        """Test hook: (nodes uint32[n_nodes, 20], primitive index per leaf-order slot, info, bounds) of the resident tree."""
        info = np.zeros(4, dtype=np.uint64)
        bounds = np.zeros(7, dtype=np.float32)
        self._ck(self.lib.crt_read_bvh(self.h, None, 0, None, 0, info.ctypes.data, bounds.ctypes.data))
        nodes = np.zeros((int(info[0]), int(info[3])), dtype=np.uint32)
        idx = np.zeros(int(info[1]), dtype=np.uint32)
        self._ck(self.lib.crt_read_bvh(self.h, nodes.ctypes.data, nodes.size, idx.ctypes.data, idx.size, info.ctypes.data, bounds.ctypes.data))
        return nodes, idx, info, bounds

    def test_detmath(self, x: np.ndarray):
        x = np.ascontiguousarray(x, dtype=np.float32)
        outs = [np.empty_like(x) for _ in range(4)]
        self._ck(self.lib.crt_test_detmath(self.h, x.ctypes.data, x.size, *[o.ctypes.data for o in outs]))
        return outs
