#!/bin/bash
# round 2 (gpurun --gpus 8; r02k = the r02f list again at HEAD, with the chunk-sized pool, plus N = 4 and N = 1 on the same box): strong scaling with the collective in every step on 8 B200s -- config 3 (split-spp and tiles), config 4
# (4K, 1024 spp per step), config 5 -- and the single-process group over 8 devices
O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/r02k_smi.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29621"
timeout 600 $TR bench.py --gpus 8 --steps 8 --warmup 3 > $O/r02k_bench_n8_strong_spp.json 2> $O/r02k_bench_n8_strong_spp.err
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29623"
timeout 600 $TR4 bench.py --gpus 4 --steps 8 --warmup 3 > $O/r02k_bench_n4_strong_spp.json 2> $O/r02k_bench_n4_strong_spp.err
timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-also > $O/r02k_bench_n1.json 2> $O/r02k_bench_n1.err
timeout 600 $TR bench.py --gpus 8 --steps 8 --warmup 3 --shard tiles > $O/r02k_bench_n8_strong_tiles.json 2> $O/r02k_bench_n8_strong_tiles.err
timeout 900 $TR bench.py --gpus 8 --steps 4 --warmup 3 --width 3840 --height 2160 --spp-per-step 1024 > $O/r02k_bench_n8_cfg4_4k_1024spp.json 2> $O/r02k_bench_n8_cfg4_4k_1024spp.err
timeout 900 $TR bench.py --gpus 8 --steps 4 --warmup 3 --workload soup10m --spp-per-step 128 > $O/r02k_bench_n8_soup10m_128spp.json 2> $O/r02k_bench_n8_soup10m_128spp.err
timeout 600 python - > $O/r02k_group8.txt 2>&1 <<'PY'
import time, numpy as np
from computeraytracer_b200 import scenes
from computeraytracer_b200.renderer import Renderer, SHARD_SPLIT_SPP, SHARD_TILES
ps = scenes.mesh_scene(1920, 1080)
for mode, name in ((SHARD_SPLIT_SPP, "split-spp"), (SHARD_TILES, "tiles")):
    with Renderer(devices=list(range(8)), shard_mode=mode) as g:
        g.set_scene(ps); g.reset(); g.reset_counters()
        g.frame(256); fb = g.framebuffer()          # warm-up: uploads, pools, graphs
        g.reset(); g.reset_counters()
        t0 = time.perf_counter()
        for _ in range(4):
            g.frame(256); fb = g.framebuffer()
        dt = time.perf_counter() - t0
        c = g.counters(); ms, nb = g.last_combine()
        print(f"single-process group of 8, {name}: {c['rays_ext'] / dt / 1e6:.0f} Mrays/s e2e ({dt / 4 * 1e3:.1f} ms per 256-spp step incl. readback), fused peer combine {ms:.3f} ms, {nb / 1e6:.1f} MB over NVLink", flush=True)
PY
