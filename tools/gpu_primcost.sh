#!/bin/bash
# Tuning run (GPU box): SAH triangle weight of the Wide8 collapse on the two scale scenes.
cd "$(dirname "$0")/.."
for c in 0.3 0.5 0.75 1.0 1.5; do
  for w in mesh1m soup10m; do
    CRT_BVH_PRIM_COST=$c python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-also --workload $w 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); r = d['roofline']
print('primcost $c $w', round(d['value']), 'Mrays/s', round(d['ms_per_step'], 1), 'ms nodes/ray', round(r['nodes_per_ray'], 2), 'tris/ray', round(r['tris_per_ray'], 2), 'bvh_nodes', d['config']['bvh_nodes'], {a: round(b, 1) for a, b in r['kernels_ms_per_step'].items()})
"
  done
done
