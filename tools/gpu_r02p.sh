#!/bin/bash
# round 2: does AUTO's "single frames of linear-list scenes -> megakernel" rule (r01r) survive wf_tail?  Cornell 1080p, single-frame dispatches
O=gpurun_out; mkdir -p $O
python - > $O/r02p_single_frame_cornell.txt 2>&1 <<'PY'
import time, torch
from computeraytracer_b200 import scenes
from computeraytracer_b200.renderer import Renderer, default_params
ps = scenes.default_scene(1920, 1080)
with Renderer(0) as r:
    st = torch.cuda.Stream(0); r.set_stream(st.cuda_stream)
    r.set_scene(ps)
    for name, kw in (("megakernel", dict(kernel=1)), ("wavefront", dict(kernel=2)), ("wavefront, tail at 2^20", dict(kernel=2, wf_tuning=21 << 8)),
                     ("wavefront, tail at 2^22", dict(kernel=2, wf_tuning=23 << 8)), ("wavefront, no tail", dict(kernel=2, wf_tuning=31 << 8))):
        for n in (1, 2, 4):
            r.set_params(default_params(**kw)); r.reset(); r.reset_counters()
            for _ in range(8): r.frame(n)
            r.synchronize(); c0 = r.counters()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(64): r.frame(n)
            e1.record(st); st.synchronize()
            c1 = r.counters(); ms = e0.elapsed_time(e1)
            print(f"{name:28s} frames/dispatch {n}: {ms / 64:.3f} ms per dispatch, {(c1['rays_ext'] - c0['rays_ext']) / ms / 1e3:.0f} Mrays/s", flush=True)
PY
