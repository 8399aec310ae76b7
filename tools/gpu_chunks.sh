#!/bin/bash
# Tuning run (GPU box): frames per chunk and path-slot pool on the headline workload.
cd "$(dirname "$0")/.."
B="python bench.py --no-cpu-baseline --no-also"
summ() { python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); k = d['roofline']['kernels_ms_per_step']
print('$1', round(d['value']), 'Mrays/s', round(d['ms_per_step'], 1), 'ms/step', 'e2e', round(d['e2e']['value']), {a: round(b, 1) for a, b in k.items()}, 'launches', d['gpu_launches'], 'frac', round(d['roofline']['frac'], 3))
"; }
$B --steps 4 --warmup 3 | summ default_256spp_chunk256_pool64M
$B --steps 4 --warmup 3 --spp-per-step 64 --frames-per-launch 64 --pool 33554432 | summ spp64_chunk64_pool32M
$B --steps 4 --warmup 3 --spp-per-step 256 --frames-per-launch 256 --pool 33554432 | summ spp256_chunk256_pool32M
$B --steps 4 --warmup 3 --spp-per-step 256 --frames-per-launch 256 --pool 100663296 | summ spp256_chunk256_pool96M
$B --steps 3 --warmup 2 --workload soup10m | summ soup10m_16spp
$B --steps 3 --warmup 2 --workload soup10m --spp-per-step 64 | summ soup10m_64spp
