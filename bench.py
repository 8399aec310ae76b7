#!/usr/bin/env python
"""bench.py -- throughput of the path-tracing hot path on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload ...]

One JSON line on stdout (rank 0).  A *step* is one batch of `--spp-per-step` frames (one sample per
pixel each) of the workload; `value` is primary+bounce rays per second (intersect() calls from
path_trace, ComputeShader.wgsl:135) with all inputs resident in HBM; `e2e` is the same metric through
the public API with the per-step host<->device copies inside the timed region.

Workloads (BASELINE.json configs):
  mesh1m   config 3: Cornell walls + light and a 1,048,576-triangle displaced sphere, 1920x1080   (default)
           (with --width 3840 --height 2160 --spp-per-step 1024: config 4)
  cornell  config 2: the reference's default scene at 1920x1080
  soup10m  config 5: 10,000,000 random triangles, max_depth 8

N > 1 (torchrun, one rank per GPU): STRONG scaling, as BASELINE.json states the configs -- the step's spp are split over
the ranks (split-spp: rank r renders the frames with (sample-1) % N == r; `--shard tiles` splits the image instead), and
every timed step ends with the collective: crt_combine (ncclReduce of the partial sums, or the gather of the owned tiles,
inside libcrt) and rank 0's tone-map.  `--scaling weak` gives every rank the full spp per step (round-1 form).
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
FLOP_PER_RAY_LINEAR = 960.0   # ~0.96 kFLOP per intersect() on the reference's default scene, as written in the WGSL

# frames per step = the samples per pixel BASELINE.json quotes for the config (configs[2], [1], [4])
CONFIG_SPP = {"mesh1m": 256, "cornell": 64, "soup10m": 16}
TRAFFIC_FILE = os.path.join(ROOT, "profiles", "r02_traffic.json")   # ncu dram__bytes per ray per kernel, rewritten by tools/gpu_round.sh


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="mesh1m", choices=["mesh1m", "cornell", "soup10m"])
    ap.add_argument("--spp-per-step", type=int, default=0, help="frames (samples per pixel) per step, all ranks together; 0 = the spp BASELINE.json quotes for the workload")
    ap.add_argument("--shard", default="spp", choices=["spp", "tiles"])
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"], help="N > 1: split the step's spp over the ranks (strong) or give every rank the full spp (weak)")
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--frames-per-launch", type=int, default=0)
    ap.add_argument("--pool", type=int, default=0, help="wavefront path slots in flight (0 = library default)")
    ap.add_argument("--kernel", default="auto", choices=["auto", "mega", "wavefront"], help="force one form of the hot path (default: chosen from the scene)")
    ap.add_argument("--wf-split", action="store_true", help="wavefront: wf_shade + wf_regen kernels instead of the single wf_logic")
    ap.add_argument("--bvh", default="auto", choices=["auto", "host", "device"], help="BVH builder: host binned SAH or device PLOC (auto: library default)")
    ap.add_argument("--wf-tuning", type=lambda s: int(s, 0), default=0, help="crt_params.wf_tuning bits (include/crt.h)")
    ap.add_argument("--bvh-tuning", type=lambda s: int(s, 0), default=0, help="crt_params.bvh_tuning bits: bins | 0x200 exact slots")
    ap.add_argument("--bvh-prim-cost", type=float, default=0.0)
    ap.add_argument("--max-depth", type=int, default=-1, help="override the workload's max_depth (experiments: 0 = primary rays only)")
    ap.add_argument("--no-l2-flush", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-also", action="store_true", help="skip the secondary measurements (Cornell config 2, single-frame dispatches, FP32 peak)")
    return ap.parse_args()


def workload_scene(name: str, width: int, height: int):
    from computeraytracer_b200 import scenes
    if name == "mesh1m":
        return scenes.mesh_scene(width, height), 100, "config 3: 1,048,576-triangle displaced sphere + Cornell walls/light"
    if name == "cornell":
        return scenes.default_scene(width, height), 100, "config 2: reference default scene (16 patches, 2 spheres)"
    return scenes.soup_scene(width, height), 8, "config 5: 10,000,000-triangle random soup + Cornell walls/light, max_depth 8"


def config_of(args, spp, max_depth, desc, world):
    """The `config` object, the same in both arms (the reference arm states its bounded sample in `cpu_baseline.sample`)."""
    return {"workload": f"{args.workload} ({desc})", "resolution": f"{args.width}x{args.height}", "spp_per_step": spp,
            "max_depth": max_depth, "n_gpus": world,
            "shard": "none" if world == 1 else f"{args.shard}, {args.scaling}",
            "l2": "not flushed" if args.no_l2_flush else "flushed between steps (256 MiB write)"}


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
def oracle_scene(ps, max_depth):
    """The CPU oracle on the SAME scene and resolution; the bounded sample is a number of whole frames."""
    from tests import orc   # bench.py's cpu_baseline / reference legs are allowed to use the oracle
    n_tri = int((ps.primitives["category"] == 2).sum())
    return orc.load().scene(ps, orc.default_params(max_depth=max_depth), use_bvh=n_tri > 0)


def run_reference(args):
    """--impl reference: the reference's own algorithm on the host cores (the CPU oracle port of its WGSL; the
    WGSL itself cannot run here: no graphics stack, no node).  Same scene, resolution and max_depth as the B200 arm; one
    step = one frame (1 spp over the full image) -- a bounded sample of the step's spp, which a rate metric allows."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    ps, max_depth, desc = workload_scene(args.workload, args.width, args.height)
    spp = args.spp_per_step or CONFIG_SPP[args.workload]
    threads = os.cpu_count() or 1
    sc = oracle_scene(ps, max_depth)
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
    sample_desc = f"1 frame (1 spp over the full {args.width}x{args.height} image) per step instead of {spp}, {threads} threads over image rows"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak" if world == 1 else args.scaling,
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_of(args, spp, max_depth, desc, max(world, args.gpus)),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample_desc},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "total_mrays_s": (rays + shadow) / dt / 1e6, "gpu_launches": 0,
        "note": "reference WGSL via a CPU WebGPU adapter is unavailable on this box; this is the CPU port (oracle/) of the same algorithm",
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ B200 arm
def time_steps(r, stream, torch, steps, frames, flush_buf, combine):
    """K steps, each bracketed by its own CUDA events on the renderer's stream; L2 flushed between steps.  A step is the
    frames plus -- on more than one GPU -- the combine (collective + rank 0's tone-map), all enqueued asynchronously: the host
    does not wait between steps (the combine is timed by events of its own on the same stream), so the K steps are one
    stream of work between the caller's barriers, as the contract states it."""
    evs = []
    cevs = []
    for _ in range(steps):
        if flush_buf is not None:
            with torch.cuda.stream(stream):
                flush_buf.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        r.frame(frames)
        if combine:
            c0 = torch.cuda.Event(enable_timing=True)
            c0.record(stream)
            r.combine(0)
        e1.record(stream)
        evs.append((e0, e1))
        if combine:
            cevs.append((c0, e1))
    stream.synchronize()
    comb = []
    if combine:
        nbytes = r.last_combine()[1]              # the same every step
        comb = [(a.elapsed_time(b), nbytes) for a, b in cevs]
    return [a.elapsed_time(b) for a, b in evs], comb


def run_b200(args):
    import torch
    import torch.distributed as dist
    from computeraytracer_b200 import multi
    from computeraytracer_b200.renderer import (BVH_BUILDER_DEVICE, BVH_BUILDER_HOST_SAH, SHARD_SPLIT_SPP, SHARD_TILES, Renderer)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: libcrt has no CPU fallback")
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    device = local
    torch.cuda.set_device(device)
    stream = torch.cuda.Stream(device)
    dev = f"cuda:{device}"

    ps, max_depth, desc = workload_scene(args.workload, args.width, args.height)
    if args.max_depth >= 0:
        max_depth = args.max_depth
    r = Renderer(device)
    r.set_stream(stream.cuda_stream)
    bvh_bits = args.bvh_tuning | {"auto": 0, "host": BVH_BUILDER_HOST_SAH, "device": BVH_BUILDER_DEVICE}[args.bvh]
    base = dict(max_depth=max_depth, frames_per_launch=args.frames_per_launch, wf_slots=args.pool,
                kernel={"auto": 0, "mega": 1, "wavefront": 2}[args.kernel], wf_tuning=args.wf_tuning, bvh_tuning=bvh_bits,
                bvh_prim_cost=args.bvh_prim_cost, flags=4 if args.wf_split else 0)
    r.set_params(**base)
    t0 = time.perf_counter()
    r.set_scene(ps)
    st = r.scene_stats()   # forces the scene-upload module to run (SoA repack + BVH build)
    r.synchronize()
    setup_s = time.perf_counter() - t0
    # a scene edit: binding 8 set again -> upload + tree once more (the first build of a process also pays CUDA's lazy module
    # loading of the builder's kernels; this one does not)
    scene_edit = None
    if st["n_triangles"] and rank == 0 and not args.no_also:
        t1 = time.perf_counter()
        r.set_binding(8, ps.bindings()[8])
        st_e = r.scene_stats(); r.synchronize()
        scene_edit = {"bvh_build_ms": st_e["tree_us"] / 1e3, "scene_upload_ms": st_e["build_us"] / 1e3, "wall_s": time.perf_counter() - t1,
                      "builder": st_e["builder"], "note": "crt_set_binding(8) with the same primitives + the lazy re-upload: what a scene edit costs"}

    spp = args.spp_per_step or CONFIG_SPP[args.workload]
    combine = world > 1
    if world > 1:
        mode = SHARD_SPLIT_SPP if args.shard == "spp" else SHARD_TILES
        multi.init(r, mode)            # NCCL communicator inside libcrt; torch.distributed only carried the 128-byte id
        frames = spp * world if (args.scaling == "weak" and args.shard == "spp") else spp
        scaling = args.scaling if args.shard == "spp" else "strong"
    else:
        frames, scaling = spp, "weak"
    flush_buf = None if args.no_l2_flush else torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    # ---- calibration pass with the detailed counters (separate, slower kernel variant; not timed)
    r.reset(); r.reset_counters(); r.enable_stats(True)
    r.frame(frames); r.synchronize()
    cal = r.counters()
    r.enable_stats(False)
    cal_rays = cal["rays_ext"] + cal["rays_shadow"]
    bytes_per_ray = (NODE_BYTES * cal["node_visits"] + TRI_BYTES * cal["tri_tests"] + QUEUE_RAY_BYTES * cal["bvh_rays"]) / max(cal_rays, 1)
    # wf_logic: every extension ray is one pass over its slot's record (written when the ray is made, read when its hit is
    # consumed); every path stages 16 B of XYZ.  Per (extension + shadow) ray, like bytes_per_ray.
    logic_bytes_per_ray = (SHADE_PASS_BYTES * cal["rays_ext"] + 16 * cal["paths"]) / max(cal_rays, 1)

    # ---- instrumented pass: CUDA events between the hot-path launches (host-driven loop) give the per-kernel-class times the
    # roofline is computed from.  The headline below runs the default path (the loop on the device) and is timed as a whole.
    r.reset(); r.reset_counters()
    r.enable_kernel_times(True)
    time_steps(r, stream, torch, 1, frames, flush_buf, combine)
    r.reset_counters()
    n_inst = max(1, min(3, args.steps))
    ms_inst, _ = time_steps(r, stream, torch, n_inst, frames, flush_buf, combine)
    kt = r.kernel_times()
    d_inst = r.counters()
    r.enable_kernel_times(False)

    # ---- device-timed region, default path
    r.reset(); r.reset_counters()
    time_steps(r, stream, torch, args.warmup, frames, flush_buf, combine)    # includes the warm-up combines (NCCL channels set up)
    c0 = r.counters()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(device)
    with ClockSampler(device) as clk:
        ms, comb = time_steps(r, stream, torch, args.steps, frames, flush_buf, combine)
        torch.cuda.synchronize(device)
        if world > 1:
            dist.barrier()
    clocks = clk.summary()
    c1 = r.counters()
    d = {k: c1[k] - c0[k] for k in c1}
    total_ms = torch.tensor([sum(ms)], dtype=torch.float64, device=dev)
    tot = torch.tensor([d["rays_ext"], d["rays_shadow"], d["paths"], d["launches"]], dtype=torch.float64, device=dev)
    comb_t = torch.tensor([sum(c[0] for c in comb) / max(len(comb), 1), float(sum(c[1] for c in comb)) / max(len(comb), 1)],
                          dtype=torch.float64, device=dev)
    comb_max = comb_t.clone()
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        dist.all_reduce(comb_max, op=dist.ReduceOp.MAX)
        dist.all_reduce(comb_t, op=dist.ReduceOp.SUM)
    job_ms = float(total_ms.item())
    rays_ext, rays_shadow, paths, launches = [float(x) for x in tot.tolist()]
    value = rays_ext / (job_ms * 1e-3) / 1e6

    # ---- roofline of the dominant kernel class, this rank's launches, from the events of the instrumented pass
    peak, peak_src = measured_peaks()
    kt = {k: v for k, v in kt.items() if v[1] > 0}
    kernel = max(kt, key=lambda k: kt[k][0]) if kt else "pt_megakernel"
    k_ms, k_n = kt.get(kernel, (sum(ms_inst), max(d_inst["launches"], 1)))
    launch_ms = k_ms / max(k_n, 1)
    inst_rays = d_inst["rays_ext"] + d_inst["rays_shadow"]
    pixels_per_step = args.width * args.height / (world if (world > 1 and args.shard == "tiles") else 1)
    l2_resident = (st["bvh_bytes"] + st["tri_bytes"]) < 100e6
    if kernel == "pt_megakernel":
        total_bytes = bytes_per_ray * inst_rays + PIXEL_BYTES * pixels_per_step * k_n
        what = "80 B x BVH nodes visited + 48 B x triangles tested (device counters) + 36 B x pixels per launch"
    elif kernel in ("wf_shade", "wf_logic"):   # path-state traffic: S0..S5 + H read and written once per bounce, queue entries
        total_bytes = logic_bytes_per_ray * inst_rays
        what = "232 B of path state (S0..S5, H read + written, live-list entry) x extension rays + 16 B of staged XYZ x paths (device counters)"
    else:   # wf_traverse (a BVH scene): the traversal kernel's own fetches
        total_bytes = bytes_per_ray * inst_rays
        what = "80 B x BVH nodes visited + 48 B x triangles tested + 84 B x rays queued (device counters of a calibration step)"
    algo_bytes = total_bytes / max(k_n, 1)
    achieved = algo_bytes / (launch_ms * 1e-3) / 1e9
    if st["n_triangles"] == 0:
        regime = "linear list: FP32 pipe / instruction issue (HBM sees 36 B per pixel and launch + the wavefront's path state)"
    elif l2_resident:
        regime = "L2-resident scene: fetch bandwidth out of L2/L1 and load latency x lane occupancy; `frac` is a fraction of the HBM copy peak used as a yardstick, `dram_frac` is the real DRAM utilisation"
    else:
        regime = "scene larger than L2: HBM bandwidth"
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "dram_frac": None, "regime": regime,
                "kernel": kernel, "launch_ms": launch_ms, "launches": k_n, "share_of_step": k_ms / max(sum(ms_inst), 1e-9),
                "algorithmic_bytes_per_launch": algo_bytes,
                "bytes_per_ray": bytes_per_ray, "nodes_per_ray": cal["node_visits"] / max(cal_rays, 1),
                "tris_per_ray": cal["tri_tests"] / max(cal_rays, 1), "bvh_rays_per_ray": cal["bvh_rays"] / max(cal_rays, 1),
                "peak_source": peak_src,
                "kernels_ms_per_step": {k: v[0] / n_inst for k, v in kt.items()},
                "instrumented_ms_per_step": sum(ms_inst) / n_inst,
                "note": "algorithmic bytes = " + what + "; fetches served by L1/L2 count, so they exceed DRAM traffic "
                        "when the scene is L2-resident (see DESIGN.md section 6).  Per-kernel times: CUDA events between the launches of "
                        "an instrumented pass (host-driven loop); the headline runs the device-side loop and is timed as a whole"}
    # both wavefront kernels side by side (they take about half of a step each, so which one is "dominant" can flip
    # from run to run): wf_traverse against its fetch bytes, wf_logic against its path-state bytes
    traffic = {}
    if os.path.exists(TRAFFIC_FILE):   # ncu-measured DRAM bytes per ray per kernel at this HEAD (tools/gpu_round.sh rewrites the file)
        try:
            with open(TRAFFIC_FILE) as f:
                traffic = json.load(f).get(args.workload, {})
        except Exception:
            traffic = {}
    if "wf_traverse" in kt and "wf_logic" in kt:
        both = {}
        for name, tb in (("wf_traverse", bytes_per_ray * inst_rays), ("wf_logic", logic_bytes_per_ray * inst_rays)):
            ms_k, n_k = kt[name]
            gbs = tb / max(ms_k * 1e-3, 1e-12) / 1e9
            both[name] = {"achieved": gbs, "frac": gbs / peak, "launch_ms": ms_k / max(n_k, 1), "launches": n_k,
                          "share_of_step": ms_k / max(sum(ms_inst), 1e-9)}
            per_ray = traffic.get(name, {}).get("dram_bytes_per_ray")
            if per_ray is not None and args.width == 1920 and args.height == 1080:
                dram = per_ray * inst_rays
                both[name]["dram_gbs"] = dram / max(ms_k * 1e-3, 1e-12) / 1e9
                both[name]["dram_frac"] = both[name]["dram_gbs"] / peak
                both[name]["traffic_over_algorithmic"] = dram / max(tb, 1)
        roofline["per_kernel"] = both
    per_ray = traffic.get(kernel, {}).get("dram_bytes_per_ray")
    if per_ray is not None and args.width == 1920 and args.height == 1080:
        roofline["traffic"] = per_ray * inst_rays / max(k_n, 1)
        roofline["dram_frac"] = roofline["traffic"] / (launch_ms * 1e-3) / 1e9 / peak
        roofline["traffic_source"] = os.path.relpath(TRAFFIC_FILE, ROOT) + " (ncu dram__bytes_read.sum + dram__bytes_write.sum per launch, cold L2 per launch under ncu)"

    # ---- end to end through the public API with host buffers
    fb_host = torch.empty((args.height, args.width, 4), dtype=torch.uint8).pin_memory().numpy()
    cam_host = ps.camera.copy()
    r.reset(); r.reset_counters()
    r.frame(frames)
    if combine:
        r.combine(0)
    r.framebuffer(fb_host)   # warm-up
    e0 = r.counters()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(device)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r.set_camera(cam_host)          # H2D: the step's input (64 B camera; the scene stays resident)
        r.frame(frames)
        if combine:
            r.combine(0)                # the collective, then the image is rank 0's
        if rank == 0:
            r.framebuffer(fb_host)      # D2H: the step's result (RGBA8 frame, combined), synchronises
        else:
            r.synchronize()
    r.synchronize()
    e2e_s = time.perf_counter() - t0
    e1 = r.counters()
    e2e_t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    e2e_r = torch.tensor([e1["rays_ext"] - e0["rays_ext"]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
        dist.all_reduce(e2e_r, op=dist.ReduceOp.SUM)
    e2e_val = float(e2e_r.item()) / float(e2e_t.item()) / 1e6
    e2e = {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": 64,
           "d2h_bytes_per_step": int(fb_host.nbytes), "wall_s": float(e2e_t.item()),
           "setup_s_not_included": setup_s,
           "with_setup": {"value": float(e2e_r.item()) / (float(e2e_t.item()) + setup_s) / 1e6, "unit": UNIT,
                          "note": f"the same {args.steps} steps plus the one-off scene upload (SoA repack + BVH build, {setup_s:.2f} s)"},
           "note": "per step: camera upload, frames, (N > 1: crt_combine), RGBA8 framebuffer readback to pinned host memory on rank 0; "
                   "scene upload + BVH build is one-off setup, reported separately and in with_setup"}

    line = None
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": job_ms / args.steps, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": config_of(args, spp, max_depth, desc, world),
            "scene": {"primitives": st["n_prims"], "triangles": st["n_triangles"], "bvh_nodes": st["bvh_nodes"],
                      "bvh_depth": st["bvh_depth"], "scene_bytes": st["bvh_bytes"] + st["tri_bytes"], "bvh_builder": st["builder"]},
            "frames_per_step_all_ranks": frames, "total_spp": frames * args.steps,
            "ms_per_frame": job_ms / (args.steps * frames),
            "total_mrays_s": (rays_ext + rays_shadow) / (job_ms * 1e-3) / 1e6,
            "rays": {"ext": rays_ext, "shadow": rays_shadow, "paths": paths},
            "setup_s": setup_s, "scene_upload_ms": st["build_us"] / 1e3, "bvh_build_ms": st["tree_us"] / 1e3, "scene_edit": scene_edit,
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "e2e": e2e,
        }
        if combine:
            line["combine"] = {"ms_per_step_max_over_ranks": float(comb_max[0].item()), "ms_per_step_mean_over_ranks": float(comb_t[0].item()) / world,
                               "bytes_sent_per_step_all_ranks": float(comb_t[1].item()),
                               "what": ("ncclReduce(SUM) of the W*H*4 f32 partial sums into rank 0's combined buffer" if args.shard == "spp"
                                        else "grouped ncclSend/ncclRecv of the owned tiles only") + " + rank 0's tone-map, inside every timed step (libcrt crt_combine)",
                               "note": "device time between the events around crt_combine on each rank's stream; it includes waiting for the slowest rank's frames"}
            line["combine_ms"] = float(comb_max[0].item())

    # ---- rank 0, N = 1 only: CPU baseline, single-frame dispatches, FP32 peak and the secondary Cornell figure
    if rank == 0 and world == 1:
        if not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            sc = oracle_scene(ps, max_depth)
            budget_s, nf, ext, shadow, sample = 12.0, 0, 0, 0, 0
            t0 = time.perf_counter()
            while time.perf_counter() - t0 < budget_s:   # whole frames of the same scene at the same resolution until ~12 s of CPU work
                _, _, cnt = sc.render(1, sample_start=sample, threads=threads, want_fb=False)
                sample += 1; nf += 1
                ext += int(cnt[1]); shadow += int(cnt[2])
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": ext / dt / 1e6, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": f"{nf} frames (1 spp each over the full {args.width}x{args.height} image) of the same scene, same max_depth "
                                              f"({dt:.1f} s of CPU work, {threads} threads over image rows)",
                                    "total_mrays_s": (ext + shadow) / dt / 1e6}
        if not args.no_also:
            # the reference's own usage: frame() = one dispatch of one frame per refresh (src/main.js:584-621)
            def single_frames(n=64):
                r.reset(); r.reset_counters()
                for _ in range(8):
                    r.frame(1)
                r.synchronize()
                c_a = r.counters()
                ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t_a = time.perf_counter()
                ea.record(stream)
                for _ in range(n):
                    r.frame(1)
                eb.record(stream)
                t_enq = time.perf_counter() - t_a
                stream.synchronize()
                c_b = r.counters()
                msd = ea.elapsed_time(eb)
                return {"dispatches": n, "ms_per_dispatch": msd / n, "host_ms_per_call": t_enq / n * 1e3,
                        "value": (c_b["rays_ext"] - c_a["rays_ext"]) / (msd * 1e-3) / 1e6, "unit": UNIT,
                        "note": "crt_dispatch_frames(1) x n back to back (the reference's frame loop), device-timed as a whole; "
                                "host_ms_per_call is how long a call keeps the host"}
            line["single_frame"] = {args.workload: single_frames()}
            try:
                fp32 = r.measure_fp32_tflops()
            except Exception:
                fp32 = None
            line["fp32_peak_tflops_measured"] = fp32
            if st["n_triangles"] and args.bvh == "auto":
                # the other builder on the same scene: the host binned-SAH tree is the quality option (crt_params.bvh_tuning), the
                # device PLOC tree (AUTO's choice at this size) is the one that builds in milliseconds
                other = BVH_BUILDER_HOST_SAH if st["builder"] == "device_ploc" else BVH_BUILDER_DEVICE
                r.set_params(bvh_tuning=args.bvh_tuning | other)
                t1 = time.perf_counter()
                st2 = r.scene_stats(); r.synchronize()
                setup2 = time.perf_counter() - t1
                r.reset(); r.reset_counters()
                time_steps(r, stream, torch, 2, frames, flush_buf, False)
                ca = r.counters()
                ms_o, _ = time_steps(r, stream, torch, 3, frames, flush_buf, False)
                cb = r.counters()
                line["other_bvh_builder"] = {"builder": st2["builder"], "bvh_build_ms": st2["tree_us"] / 1e3, "scene_upload_s": setup2,
                                             "ms_per_step": sum(ms_o) / 3, "value": (cb["rays_ext"] - ca["rays_ext"]) / (sum(ms_o) * 1e-3) / 1e6,
                                             "unit": UNIT, "bvh_nodes": st2["bvh_nodes"], "bvh_depth": st2["bvh_depth"]}
                r.set_params(bvh_tuning=bvh_bits)
            if args.workload != "cornell":
                from computeraytracer_b200 import scenes
                r.set_scene(scenes.default_scene(args.width, args.height))
                r.set_params(max_depth=100)
                r.reset(); r.reset_counters()
                spp2 = CONFIG_SPP["cornell"]
                time_steps(r, stream, torch, 2, spp2, flush_buf, False)
                c0 = r.counters()
                ms2, _ = time_steps(r, stream, torch, 8, spp2, flush_buf, False)
                c1 = r.counters()
                tot2 = (c1["rays_ext"] + c1["rays_shadow"] - c0["rays_ext"] - c0["rays_shadow"])
                also = {"workload": "cornell (config 2: reference default scene, 1920x1080, max_depth 100)",
                        "value": (c1["rays_ext"] - c0["rays_ext"]) / (sum(ms2) * 1e-3) / 1e6, "unit": UNIT,
                        "total_mrays_s": tot2 / (sum(ms2) * 1e-3) / 1e6,
                        "ms_per_frame": sum(ms2) / (8 * spp2), "steps": 8, "spp_per_step": spp2}
                if fp32:
                    tf = tot2 * FLOP_PER_RAY_LINEAR / (sum(ms2) * 1e-3) / 1e12
                    also["roofline"] = {"bound": "fp32", "achieved": tf, "peak": fp32, "unit": "TFLOP/s", "frac": tf / fp32,
                                        "note": "0.96 kFLOP per intersect() as written in the WGSL (SURVEY 8d) x all rays, against the FP32 FMA rate "
                                                "measured in this run; the kernels hoist ray invariants, so the executed count is about half"}
                line["also"] = also
                line["single_frame"]["cornell"] = single_frames()
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
