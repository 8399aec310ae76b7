#!/usr/bin/env python
"""Builder evaluation without a GPU: builds the Wide8 tree of a scale scene with the library's own builder
(crt_host_build_bvh) and walks it on the CPU (tools/wide8_sim.c, wf_traverse's visiting order) with a path-traced ray
mix -- primary rays, cosine-distributed bounce rays from the hits on the mesh and on the room's walls, bounded any-hit
shadow rays towards the light -- and prints BVH nodes fetched and triangles tested per queued ray.  Tree shape never
changes results (DESIGN.md "Conservative culling"), only these counts, which is what wf_traverse's time follows.
On mesh1m it gives 8.4 nodes / 2.6 triangles per queued ray where the device counters say 9.4 / 3.1.

    python tools/bvh_eval.py [mesh|soup] [n_lon or n_triangles] [KEY=VALUE ...]
        BVH_TUNING=<crt_params.bvh_tuning>   e.g. 128 (bins) | 0x200 (exact slots) | 0x800 (the device build's PLOC algorithm)
        BVH_PRIM_COST=<float>
    SIM_ORDER=1 python tools/bvh_eval.py ...     children visited by entry distance instead of by ray octant
"""
import ctypes as C
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from computeraytracer_b200 import scenes  # noqa: E402
from computeraytracer_b200.build import build_libcrt  # noqa: E402

ROOM = 555.0
ORDER = int(os.environ.get("SIM_ORDER", "0"))   # 1: children by entry distance with culling on the way back (bound for any child ordering)
TMIN = 1e-3


def load_sim():
    so = os.path.join(tempfile.gettempdir(), "wide8_sim.so")
    src = os.path.join(ROOT, "tools", "wide8_sim.c")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.run(["gcc", "-O2", "-shared", "-fPIC", "-o", so, src, "-lm"], check=True)
    lib = C.CDLL(so)
    lib.wide8_trace.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    lib.wide8_trace.restype = None
    lib.wide8_warp_sim.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.wide8_warp_sim.restype = None
    return lib


def build_tree(prims):
    lib = C.CDLL(os.environ.get("BVH_LIB") or build_libcrt())   # BVH_LIB: a host-only build of the builder (experiments)
    f = lib.crt_host_build_bvh
    f.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_uint32, C.c_float]
    tuning = int(os.environ.get("BVH_TUNING", "0"), 0)
    prim_cost = float(os.environ.get("BVH_PRIM_COST", "0"))
    raw = np.ascontiguousarray(prims).view(np.uint8)
    info = np.zeros(4, dtype=np.uint64); bounds = np.zeros(7, dtype=np.float32)
    t0 = time.perf_counter()
    assert f(raw.ctypes.data, raw.size, os.cpu_count() or 1, None, 0, None, 0, info.ctypes.data, bounds.ctypes.data, tuning, prim_cost) == 0
    dt = time.perf_counter() - t0
    nodes = np.zeros(int(info[0]) * int(info[3]), dtype=np.uint32); src = np.zeros(int(info[1]), dtype=np.uint32)
    assert f(raw.ctypes.data, raw.size, os.cpu_count() or 1, nodes.ctypes.data, nodes.size, src.ctypes.data, src.size, info.ctypes.data, bounds.ctypes.data, tuning, prim_cost) == 0
    return nodes, src, info, bounds, dt


def slab(o, d, lo, hi, tmax):
    """(enters within [TMIN, tmax], exit distance) of rays against the box [lo, hi]"""
    with np.errstate(divide="ignore", invalid="ignore"):
        t0 = (lo - o) / d; t1 = (hi - o) / d
    a = np.minimum(t0, t1); b = np.maximum(t0, t1)
    a = np.where(np.isnan(a), -np.inf, a); b = np.where(np.isnan(b), np.inf, b)
    tn = np.maximum(a.max(axis=1), TMIN); tf = b.min(axis=1)
    return tn <= np.minimum(tf, tmax), tf


def cosine_dirs(n, rng):
    u, v = rng.random(n.shape[0]), rng.random(n.shape[0])
    r, th = np.sqrt(u), 2 * np.pi * v
    x, y, z = r * np.cos(th), r * np.sin(th), np.sqrt(np.maximum(0, 1 - u))
    up = np.where((np.abs(n[:, 2]) < 0.999)[:, None], np.array([0, 0, 1.0]), np.array([1.0, 0, 0]))
    t = np.cross(up, n); t /= np.linalg.norm(t, axis=1, keepdims=True)
    b = np.cross(n, t)
    return t * x[:, None] + b * y[:, None] + n * z[:, None]


def main():
    kind = sys.argv[1] if len(sys.argv) > 1 and "=" not in sys.argv[1] else "mesh"
    size = None
    for a in sys.argv[2:]:
        if "=" in a:
            k, v = a.split("=", 1); os.environ[k] = v
        else:
            size = int(a)
    if kind == "mesh":
        n_lon = size or 1024
        ps = scenes.mesh_scene(1920, 1080, n_lon=n_lon, n_lat=n_lon // 2)
    else:
        ps = scenes.soup_scene(1920, 1080, n=size or 2_000_000)
    prims = ps.primitives
    nodes, src, info, bounds, dt = build_tree(prims)
    tri = np.ascontiguousarray(np.concatenate([prims["data1"][src], prims["data2"][src], prims["data3"][src]], axis=1).astype(np.float32))
    sim = load_sim()
    rng = np.random.default_rng(7)

    queued = {False: [], True: []}   # the rays as the kernel's queues hold them: extension rays, shadow rays

    def trace(o, d, tmax, anyhit):
        keep = slab(o, d, bounds[0:3], bounds[3:6], tmax)[0] & (tmax > 0)   # hits_root_box: the others are never queued
        o, d, tmax = o[keep], d[keep], tmax[keep]
        rays = np.zeros((o.shape[0], 8), dtype=np.float32)
        rays[:, 0:3] = o; rays[:, 3] = tmax; rays[:, 4:7] = d; rays[:, 7] = 1.0 if anyhit else 0.0
        queued[anyhit].append(rays)
        t = np.zeros(o.shape[0], dtype=np.float32); slot = np.zeros(o.shape[0], dtype=np.int32); cnt = np.zeros((o.shape[0], 2), dtype=np.uint32)
        sim.wide8_trace(nodes.ctypes.data, int(info[0]), tri.ctypes.data, rays.ctypes.data, o.shape[0], TMIN, t.ctypes.data, slot.ctypes.data, cnt.ctypes.data, ORDER)
        return o, d, t, slot, cnt, keep

    # primary rays: every 6th pixel in x and y (camera_ray, CS:470-500, pixel centres)
    cam = ps.camera
    eye, look = cam[0:3].astype(np.float64), cam[4:7].astype(np.float64)
    W, H, fov = float(cam[11]), float(cam[12]), float(cam[13])
    w = eye - look; w /= np.linalg.norm(w)
    u = np.cross(cam[8:11].astype(np.float64), w); u /= np.linalg.norm(u)
    v = np.cross(w, u)
    vph = 2 * np.tan(fov / 2); vpw = W / H * vph
    xs, ys = np.meshgrid(np.arange(2, W, 6), np.arange(2, H, 6))
    s, t_ = (xs.ravel() + 0.5) / W, (H - ys.ravel() + 0.5) / H
    d = (-(vpw / 2) * u - (vph / 2) * v - w)[None, :] + s[:, None] * vpw * u + t_[:, None] * vph * v
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    o = np.repeat(eye[None, :], d.shape[0], axis=0)
    light_o, light_e1, light_e2 = np.array([213.0, 555.0, 227.0]), np.array([130.0, 0, 0]), np.array([0, 0, 105.0])

    rows = []
    tot_n = tot_t = tot_r = 0
    for depth in range(3):
        inside, t_exit = slab(o, d, np.zeros(3), np.full(3, ROOM), np.inf)
        tmax = np.where(inside, t_exit, 0.0).astype(np.float32)            # the walls of the room bound every ray (linear list)
        oo, dd, t, slot, cnt, keep = trace(o.astype(np.float32), d.astype(np.float32), tmax, False)
        rows.append((f"extension rays, depth {depth}", o.shape[0], oo.shape[0], cnt))
        hit = slot >= 0
        P = oo[hit].astype(np.float64) + dd[hit].astype(np.float64) * t[hit, None]
        n = np.cross(tri[slot[hit], 3:6].astype(np.float64), tri[slot[hit], 6:9].astype(np.float64))
        ln = np.linalg.norm(n, axis=1, keepdims=True); ok = ln[:, 0] > 0
        P, n, inc = P[ok], n[ok] / np.where(ln[ok] > 0, ln[ok], 1.0), dd[hit][ok].astype(np.float64)
        n = np.where((np.sum(n * inc, axis=1) > 0)[:, None], -n, n)
        # rays that end on a wall of the room: everything that was not queued, or was queued and missed the triangles
        wall = np.ones(o.shape[0], dtype=bool)
        wall[np.nonzero(keep)[0][hit]] = False
        wall &= tmax > 0
        Pw = o[wall] + d[wall] * tmax[wall, None].astype(np.float64)
        ax = np.argmin(np.minimum(np.abs(Pw), np.abs(Pw - ROOM)), axis=1)
        ar = np.arange(Pw.shape[0])
        nw = np.zeros_like(Pw); nw[ar, ax] = np.where(Pw[ar, ax] < ROOM / 2, 1.0, -1.0)
        closed = Pw[:, 2] > 1e-3                                            # the front (z = 0) is open: those rays left the scene
        P = np.concatenate([P, Pw[closed]]); n = np.concatenate([n, nw[closed]])
        if P.shape[0] == 0:
            break
        # shadow rays towards the light (bounded, any-hit), then the next bounce
        lp = light_o + rng.random((P.shape[0], 1)) * light_e1 + rng.random((P.shape[0], 1)) * light_e2
        sd = lp - P; dist = np.linalg.norm(sd, axis=1); sd /= dist[:, None]
        front = np.sum(sd * n, axis=1) > 0
        so, _, _, _, scnt, _ = trace(P[front].astype(np.float32), sd[front].astype(np.float32), dist[front].astype(np.float32), True)
        rows.append((f"shadow rays, depth {depth}", int(front.sum()), so.shape[0], scnt))
        o, d = P, cosine_dirs(n, rng)
    print(f"{kind}: {int(info[1]):,} triangles, {int(info[0]):,} wide nodes, depth {int(info[2])}, build {dt:.2f} s"
          + "".join(f", {k}={v}" for k, v in os.environ.items() if k.startswith("BVH_")))
    print(f"{'ray class':<28}{'rays':>9}{'queued':>9}{'nodes/queued':>14}{'tris/queued':>13}{'max nodes':>11}")
    for name, n_all, n_q, cnt in rows:
        if n_q:
            print(f"{name:<28}{n_all:>9}{n_q:>9}{cnt[:, 0].mean():>14.2f}{cnt[:, 1].mean():>13.2f}{int(cnt[:, 0].max()):>11}")
            tot_n += int(cnt[:, 0].sum()); tot_t += int(cnt[:, 1].sum()); tot_r += n_q
    print(f"{'all queued rays':<28}{'':>9}{tot_r:>9}{tot_n / tot_r:>14.2f}{tot_t / tot_r:>13.2f}   80 B x nodes + 48 B x tris = {(80 * tot_n + 48 * tot_t) / tot_r:.0f} B/ray")
    if os.environ.get("SIM_WARP"):
        # warp-level model of wf_traverse's schedule (wide8_sim.c): steps and lanes per step; ~265 instructions per node
        # step and ~80 per triangle step on the GPU
        allrays = np.ascontiguousarray(np.concatenate(queued[False] + queued[True]))
        print(f"{'schedule':<44}{'node steps/ray':>15}{'lanes':>7}{'tri steps/ray':>15}{'lanes':>7}{'nodes/ray':>11}{'tris/ray':>10}{'instr/ray':>11}")
        rows = [("kernel (1 triangle, then 1 node per trip)", 0, 1, 1)] + [(f"postponed leaves, triangle step at >= {k} lanes", 1, k, 1) for k in (2, 4, 8, 12, 16, 24)]
        rows += [(f"up to {dq} postponed groups, >= {k} lanes", 1, k, dq) for dq in (2, 4) for k in (8, 16, 24)]
        ref_res = None
        for name, mode, tri_min, dq in rows:
            out = np.zeros(6, dtype=np.uint64)
            res_t = np.zeros(allrays.shape[0], dtype=np.float32); res_hit = np.zeros(allrays.shape[0], dtype=np.int32)
            sim.wide8_warp_sim(nodes.ctypes.data, int(info[0]), tri.ctypes.data, allrays.ctypes.data, allrays.shape[0], TMIN, mode, int(os.environ.get("SIM_REFILL", "8")), tri_min, dq, out.ctypes.data, res_t.ctypes.data, res_hit.ctypes.data)
            if ref_res is None:
                ref_res = (res_t.copy(), res_hit.copy())
            else:   # a schedule may change the work, never the answers: closest t of extension rays, occluded or not for shadow rays
                ext = allrays[:, 7] == 0
                assert np.array_equal(res_t[ext], ref_res[0][ext]) and np.array_equal(res_hit, ref_res[1]), name
            ns, nl, ts, tl, nv, tv = (float(x) for x in out)
            nr = allrays.shape[0]
            print(f"{name:<44}{ns / nr:>15.3f}{nl / max(ns, 1):>7.1f}{ts / nr:>15.3f}{tl / max(ts, 1):>7.1f}{nv / nr:>11.2f}{tv / nr:>10.2f}{(265 * ns + 80 * ts) / nr:>11.1f}")


if __name__ == "__main__":
    main()
