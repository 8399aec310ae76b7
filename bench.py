#!/usr/bin/env python
"""bench.py -- throughput of the path-tracing hot path on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload ...]

One JSON line on stdout (rank 0).  A *step* is one batch of `--spp-per-step` frames (one sample per
pixel each) of the workload; `value` is primary+bounce rays per second (intersect() calls from
path_trace, ComputeShader.wgsl:135) with all inputs resident in HBM; `e2e` is the same metric through
the public API with the per-step host<->device copies inside the timed region.

Workloads (BASELINE.json configs):
  mesh1m   config 3: Cornell walls + light and a 1,048,576-triangle displaced sphere, 1920x1080   (default)
  cornell  config 2: the reference's default scene at 1920x1080
  soup10m  config 5: 10,000,000 random triangles, max_depth 8

N > 1: launched by torchrun, one rank per GPU.  Default sharding is split-spp (each rank renders
`spp-per-step` of the N*spp-per-step frames of a step: fixed work per GPU, "weak"); `--shard tiles`
splits the image instead ("strong").  The only collective is the accumulator reduce at the end.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "Mrays/s (primary+bounce) at 1080p"
UNIT = "Mrays/s"
NODE_BYTES = 80   # one 8-wide BVH node fetch (5 x 16 bytes), crt_device.cuh wide_node()
TRI_BYTES = 48    # one triangle fetch (3 x float4)
PIXEL_BYTES = 36  # accumulator read 16 + write 16 + framebuffer write 4, once per pixel per launch
SHADE_PASS_BYTES = 232  # wf_logic per extension ray: S0..S5 + H written (112) and read back (112), live-list entry written + read (8)
QUEUE_RAY_BYTES = 84  # wf_traverse per queued ray: queue entry 4 + origin 16 + direction 16 + exclude word 16 + hit read 16 + hit write 16


# frames per step = the samples per pixel BASELINE.json quotes for the config (configs[2], [1], [4])
CONFIG_SPP = {"mesh1m": 256, "cornell": 64, "soup10m": 16}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="mesh1m", choices=["mesh1m", "cornell", "soup10m"])
    ap.add_argument("--spp-per-step", type=int, default=0, help="frames (samples per pixel) per step; 0 = the spp BASELINE.json quotes for the workload")
    ap.add_argument("--shard", default="spp", choices=["spp", "tiles"])
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--frames-per-launch", type=int, default=0)
    ap.add_argument("--pool", type=int, default=0, help="wavefront path slots in flight (0 = library default)")
    ap.add_argument("--kernel", default="auto", choices=["auto", "mega", "wavefront"], help="force one form of the hot path (default: chosen from the scene)")
    ap.add_argument("--wf-split", action="store_true", help="wavefront: wf_shade + wf_regen kernels instead of the single wf_logic")
    ap.add_argument("--no-l2-flush", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-also", action="store_true", help="skip the secondary Cornell (config 2) measurement")
    return ap.parse_args()


def workload_scene(name: str, width: int, height: int):
    from computeraytracer_b200 import scenes
    if name == "mesh1m":
        return scenes.mesh_scene(width, height), 100, "config 3: 1,048,576-triangle displaced sphere + Cornell walls/light"
    if name == "cornell":
        return scenes.default_scene(width, height), 100, "config 2: reference default scene (16 patches, 2 spheres)"
    return scenes.soup_scene(width, height), 8, "config 5: 10,000,000-triangle random soup + Cornell walls/light, max_depth 8"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU every 100 ms while running."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x10: "sync_boost"}

    def __init__(self, device: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._h = None
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            self._nv = pynvml
            try:
                uuid = "GPU-" + str(torch.cuda.get_device_properties(device).uuid)
                self._h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
            except Exception:
                vis = os.environ.get("CUDA_VISIBLE_DEVICES")
                idx = int(vis.split(",")[device]) if vis and vis.split(",")[device].isdigit() else device
                self._h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._h = None
        self._t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(int(self._nv.nvmlDeviceGetClockInfo(self._h, self._nv.NVML_CLOCK_SM)))
                try:
                    bits = int(self._nv.nvmlDeviceGetCurrentClocksEventReasons(self._h))
                except Exception:
                    bits = int(self._nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
                for b, name in self.REASONS.items():
                    if bits & b:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.1)

    def __enter__(self):
        if self._h is not None:
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._h is not None:
            self._t.join(timeout=2)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": int(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------ CPU (oracle) legs
def oracle_sample(ps_full, max_depth, width, height, frames, threads):
    """Times the CPU oracle on a bounded sample: the same scene at a reduced resolution (same field of view)."""
    from tests import orc   # bench.py's cpu_baseline / reference legs are allowed to use the oracle
    import dataclasses
    cam = ps_full.camera.copy()
    cam[11], cam[12] = width, height
    ps = dataclasses.replace(ps_full, camera=cam, width=width, height=height)
    n_tri = int((ps.primitives["category"] == 2).sum())
    sc = orc.load().scene(ps, orc.default_params(max_depth=max_depth), use_bvh=n_tri > 0)
    return sc, ps


def run_reference(args):
    """--impl reference: the reference's own algorithm on the host cores (the CPU oracle port of its WGSL; the
    WGSL itself cannot run here: no graphics stack, no node).  One step = one frame of a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ps, max_depth, desc = workload_scene(args.workload, args.width, args.height)
    threads = os.cpu_count() or 1
    sw, sh = args.width // 2, args.height // 2
    sc, _ = oracle_sample(ps, max_depth, sw, sh, 1, threads)
    sample = 0
    for _ in range(args.warmup):
        sc.render(1, sample_start=sample, threads=threads, want_fb=False)
        sample += 1
    rays = 0
    shadow = 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        _, _, cnt = sc.render(1, sample_start=sample, threads=threads, want_fb=False)
        sample += 1
        rays += int(cnt[1]); shadow += int(cnt[2])
    dt = time.perf_counter() - t0
    value = rays / dt / 1e6
    sample_desc = f"{sw}x{sh} x 1 frame per step of the same scene (1/4 of the pixels), {threads} threads"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload} ({desc})", "resolution": f"{args.width}x{args.height}",
                   "max_depth": max_depth, "sample": sample_desc},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample_desc},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "total_mrays_s": (rays + shadow) / dt / 1e6, "gpu_launches": 0,
        "note": "reference WGSL via a CPU WebGPU adapter is unavailable on this box; this is the CPU port (oracle/) of the same algorithm",
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ B200 arm
def time_steps(r, stream, torch, steps, step_frames, flush_buf):
    """K steps, each bracketed by its own CUDA events on the renderer's stream; L2 flushed between steps."""
    evs = []
    for _ in range(steps):
        if flush_buf is not None:
            with torch.cuda.stream(stream):
                flush_buf.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        r.frame(step_frames)
        e1.record(stream)
        evs.append((e0, e1))
    stream.synchronize()
    return [a.elapsed_time(b) for a, b in evs]


def measure(r, stream, torch, dist, args, world, step_frames, flush_buf, device, combine_fn):
    r.reset(); r.reset_counters()
    time_steps(r, stream, torch, args.warmup, step_frames, flush_buf)
    c0 = r.counters()
    k0 = r.kernel_times()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(device)
    with ClockSampler(device) as clk:
        ms = time_steps(r, stream, torch, args.steps, step_frames, flush_buf)
        comb_ms = 0.0
        if world > 1:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            combine_fn()
            e1.record(stream)
            stream.synchronize()
            comb_ms = e0.elapsed_time(e1)
        torch.cuda.synchronize(device)
        if world > 1:
            dist.barrier()
    c1 = r.counters()
    k1 = r.kernel_times()
    d = {k: c1[k] - c0[k] for k in c1}
    kt = {k: (k1[k][0] - k0[k][0], k1[k][1] - k0[k][1]) for k in k1}
    return ms, comb_ms, d, clk.summary(), kt


def run_b200(args):
    import torch
    import torch.distributed as dist
    from computeraytracer_b200 import multi
    from computeraytracer_b200.renderer import SHARD_SPLIT_SPP, SHARD_TILES, Renderer

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: libcrt has no CPU fallback")
    device = local
    torch.cuda.set_device(device)
    stream = torch.cuda.Stream(device)

    ps, max_depth, desc = workload_scene(args.workload, args.width, args.height)
    r = Renderer(device)
    r.set_stream(stream.cuda_stream)
    t0 = time.perf_counter()
    r.set_scene(ps)
    p = r.set_params(max_depth=max_depth, frames_per_launch=args.frames_per_launch)
    if args.pool or args.wf_split or args.kernel != "auto":
        p.wf_slots = args.pool
        p.kernel = {"auto": 0, "mega": 1, "wavefront": 2}[args.kernel]
        if args.wf_split:
            p.flags |= 4   # CRT_FLAG_WF_SPLIT
        r.set_params(p)
    st = r.scene_stats()   # forces the scene-upload module to run (SoA repack + BVH build)
    r.synchronize()
    setup_s = time.perf_counter() - t0

    spp = args.spp_per_step or CONFIG_SPP[args.workload]
    if world > 1 and args.shard == "spp":
        r.set_shard(rank, world, SHARD_SPLIT_SPP)
        step_frames, scaling = spp * world, "weak"
    elif world > 1:
        r.set_shard(rank, world, SHARD_TILES)
        step_frames, scaling = spp, "strong"
    else:
        step_frames, scaling = spp, "weak"
    flush_buf = None if args.no_l2_flush else torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{device}")
    combine_fn = (lambda: multi.combine(r, stream)) if world > 1 else (lambda: None)

    # ---- calibration pass with the detailed counters (separate, slower kernel variant; not timed)
    r.reset(); r.reset_counters(); r.enable_stats(True)
    r.frame(step_frames); r.synchronize()
    cal = r.counters()
    r.enable_stats(False)
    cal_rays = cal["rays_ext"] + cal["rays_shadow"]
    bytes_per_ray = (NODE_BYTES * cal["node_visits"] + TRI_BYTES * cal["tri_tests"] + QUEUE_RAY_BYTES * cal["bvh_rays"]) / max(cal_rays, 1)
    # wf_logic: every extension ray is one pass over its slot's record (written when the ray is made, read when its hit is
    # consumed); every path stages 16 B of XYZ.  Per (extension + shadow) ray, like bytes_per_ray.
    logic_bytes_per_ray = (SHADE_PASS_BYTES * cal["rays_ext"] + 16 * cal["paths"]) / max(cal_rays, 1)

    # ---- device-timed region (events between the hot-path launches give the per-kernel-class times)
    r.enable_kernel_times(True)
    ms, comb_ms, d, clocks, kt = measure(r, stream, torch, dist, args, world, step_frames, flush_buf, device, combine_fn)
    total_ms = torch.tensor([sum(ms) + comb_ms], dtype=torch.float64, device=f"cuda:{device}")
    tot = torch.tensor([d["rays_ext"], d["rays_shadow"], d["paths"], d["launches"]], dtype=torch.float64, device=f"cuda:{device}")
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    job_ms = float(total_ms.item())
    rays_ext, rays_shadow, paths, launches = [float(x) for x in tot.tolist()]
    value = rays_ext / (job_ms * 1e-3) / 1e6

    r.enable_kernel_times(False)

    # ---- roofline of the dominant kernel class, this rank's launches, from the events of the timed region
    peak, peak_src = measured_peaks()
    kt = {k: v for k, v in kt.items() if v[1] > 0}
    kernel = max(kt, key=lambda k: kt[k][0]) if kt else "pt_megakernel"
    k_ms, k_n = kt.get(kernel, (sum(ms), max(d["launches"], 1)))
    launch_ms = k_ms / max(k_n, 1)
    my_rays = d["rays_ext"] + d["rays_shadow"]
    pixels_per_step = args.width * args.height / (world if (world > 1 and args.shard == "tiles") else 1)
    if kernel == "pt_megakernel":
        total_bytes = bytes_per_ray * my_rays + PIXEL_BYTES * pixels_per_step * k_n
        what = "80 B x BVH nodes visited + 48 B x triangles tested (device counters) + 36 B x pixels per launch"
    elif kernel in ("wf_shade", "wf_logic"):   # path-state traffic: S0..S5 + H read and written once per bounce, queue entries
        total_bytes = logic_bytes_per_ray * my_rays
        what = "232 B of path state (S0..S5, H read + written, live-list entry) x extension rays + 16 B of staged XYZ x paths (device counters)"
    else:   # wf_traverse (a BVH scene): the traversal kernel's own fetches
        total_bytes = bytes_per_ray * my_rays
        what = "80 B x BVH nodes visited + 48 B x triangles tested + 84 B x rays queued (device counters of a calibration step)"
    algo_bytes = total_bytes / max(k_n, 1)
    achieved = algo_bytes / (launch_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "kernel": kernel, "launch_ms": launch_ms, "launches": k_n, "share_of_step": k_ms / max(sum(ms), 1e-9),
                "algorithmic_bytes_per_launch": algo_bytes,
                "bytes_per_ray": bytes_per_ray, "nodes_per_ray": cal["node_visits"] / max(cal_rays, 1),
                "tris_per_ray": cal["tri_tests"] / max(cal_rays, 1), "bvh_rays_per_ray": cal["bvh_rays"] / max(cal_rays, 1),
                "peak_source": peak_src,
                "kernels_ms_per_step": {k: v[0] / args.steps for k, v in kt.items()},
                "note": "algorithmic bytes = " + what + "; fetches served by L1/L2 count, so they exceed DRAM traffic "
                        "when the scene is L2-resident (see DESIGN.md section 6)"}
    # both wavefront kernels side by side (they take about half of a step each, so which one is "dominant" can flip
    # from run to run): wf_traverse against its fetch bytes, wf_logic against its path-state bytes
    if "wf_traverse" in kt and "wf_logic" in kt:
        both = {}
        for name, tb in (("wf_traverse", bytes_per_ray * my_rays),
                         ("wf_logic", logic_bytes_per_ray * my_rays)):
            ms_k, n_k = kt[name]
            gbs = tb / max(ms_k * 1e-3, 1e-12) / 1e9
            both[name] = {"achieved": gbs, "frac": gbs / peak, "launch_ms": ms_k / max(n_k, 1), "launches": n_k,
                          "share_of_step": ms_k / max(sum(ms), 1e-9)}
        roofline["per_kernel"] = both
    prof = os.path.join(ROOT, "profiles", "r01_traffic.json")
    if os.path.exists(prof):   # ncu-measured DRAM bytes per ray of this kernel (see the file), scaled to this run's launches
        try:
            with open(prof) as f:
                per_ray = json.load(f).get(args.workload, {}).get(kernel, {}).get("dram_bytes_per_ray")
            if per_ray is not None and args.width == 1920 and args.height == 1080:
                roofline["traffic"] = per_ray * my_rays / max(k_n, 1)
                roofline["traffic_source"] = "profiles/r01_traffic.json (ncu launch list, cold L2 per launch)"
        except Exception:
            pass

    # ---- end to end through the public API with host buffers
    fb_host = torch.empty((args.height, args.width, 4), dtype=torch.uint8).pin_memory().numpy()
    cam_host = ps.camera.copy()
    r.reset(); r.reset_counters()
    r.frame(step_frames); r.framebuffer(fb_host)   # warm-up
    e0 = r.counters()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(device)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r.set_camera(cam_host)          # H2D: the step's input (64 B camera; the scene stays resident)
        r.frame(step_frames)
        r.framebuffer(fb_host)          # D2H: the step's result (RGBA8 frame), synchronises
    if world > 1:
        combine_fn()
        if rank == 0:
            r.framebuffer(fb_host)
    r.synchronize()
    e2e_s = time.perf_counter() - t0
    e1 = r.counters()
    e2e_t = torch.tensor([e2e_s], dtype=torch.float64, device=f"cuda:{device}")
    e2e_r = torch.tensor([e1["rays_ext"] - e0["rays_ext"]], dtype=torch.float64, device=f"cuda:{device}")
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
        dist.all_reduce(e2e_r, op=dist.ReduceOp.SUM)
    e2e = {"value": float(e2e_r.item()) / float(e2e_t.item()) / 1e6, "unit": UNIT, "h2d_bytes_per_step": 64,
           "d2h_bytes_per_step": int(fb_host.nbytes), "wall_s": float(e2e_t.item()),
           "setup_s_not_included": setup_s,
           "note": "per step: camera upload, frames, RGBA8 framebuffer readback to pinned host memory; "
                   "scene upload + BVH build is one-off setup, reported separately"}

    line = None
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": job_ms / args.steps, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload} ({desc})", "resolution": f"{args.width}x{args.height}",
                       "spp_per_step": spp, "frames_per_step_all_ranks": step_frames, "total_spp": step_frames * args.steps,
                       "max_depth": max_depth, "shard": ("none" if world == 1 else args.shard),
                       "l2": "not flushed" if args.no_l2_flush else "flushed between steps (256 MiB write)",
                       "primitives": st["n_prims"], "triangles": st["n_triangles"], "bvh_nodes": st["bvh_nodes"],
                       "bvh_depth": st["bvh_depth"], "scene_bytes": st["bvh_bytes"] + st["tri_bytes"]},
            "ms_per_frame": job_ms / (args.steps * spp),
            "total_mrays_s": (rays_ext + rays_shadow) / (job_ms * 1e-3) / 1e6,
            "rays": {"ext": rays_ext, "shadow": rays_shadow, "paths": paths},
            "combine_ms": comb_ms, "setup_s": setup_s, "bvh_build_ms": st["build_us"] / 1e3,
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "e2e": e2e,
        }

    # ---- rank 0, N = 1 only: CPU baseline and the secondary Cornell figure
    if rank == 0 and world == 1:
        if not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            sw, sh = args.width // 2, args.height // 2
            sc, _ = oracle_sample(ps, max_depth, sw, sh, 1, threads)
            budget_s, frames, ext, shadow, sample = 12.0, 0, 0, 0, 0
            t0 = time.perf_counter()
            while time.perf_counter() - t0 < budget_s:   # whole frames of the same scene until ~12 s of CPU work
                _, _, cnt = sc.render(1, sample_start=sample, threads=threads, want_fb=False)
                sample += 1; frames += 1
                ext += int(cnt[1]); shadow += int(cnt[2])
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": ext / dt / 1e6, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": f"{sw}x{sh} x {frames} frames of the same scene, same max_depth ({dt:.1f} s of CPU work, "
                                              f"{threads} threads over image rows)",
                                    "total_mrays_s": (ext + shadow) / dt / 1e6}
        if not args.no_also and args.workload != "cornell":
            from computeraytracer_b200 import scenes
            r.set_scene(scenes.default_scene(args.width, args.height))
            r.set_params(max_depth=100)
            r.reset(); r.reset_counters()
            spp2 = CONFIG_SPP["cornell"]
            ms2 = time_steps(r, stream, torch, 2, spp2, flush_buf)
            c0 = r.counters()
            ms2 = time_steps(r, stream, torch, 8, spp2, flush_buf)
            c1 = r.counters()
            line["also"] = {"workload": "cornell (config 2: reference default scene, 1920x1080, max_depth 100)",
                            "value": (c1["rays_ext"] - c0["rays_ext"]) / (sum(ms2) * 1e-3) / 1e6, "unit": UNIT,
                            "total_mrays_s": (c1["rays_ext"] + c1["rays_shadow"] - c0["rays_ext"] - c0["rays_shadow"]) / (sum(ms2) * 1e-3) / 1e6,
                            "ms_per_frame": sum(ms2) / (8 * spp2), "steps": 8, "spp_per_step": spp2}
    r.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
