#!/bin/bash
# A/B runs on the GPU box: each line of the spec file is  name | library (path or -) | ENV=.. (or -) | bench args
# Prints one summary line per run and appends it to gpurun_out/<tag>_ab.txt.   Usage: tools/gpu_ab.sh <tag> <specfile>
TAG=$1; SPEC=$2; O=gpurun_out; mkdir -p $O
while IFS='|' read -r name lib envs args; do
  name=$(echo $name); lib=$(echo $lib); envs=$(echo $envs); [ -z "$name" ] && continue
  [ "$lib" = "-" ] && lib=computeraytracer_b200/libcrt.so
  [ "$envs" = "-" ] && envs=""
  line=$(env CRT_LIBCRT=$PWD/$lib $envs python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-also $args 2>$O/${TAG}_${name}.err | tail -1 | python -c "
import sys, json
l = sys.stdin.read().strip()
try:
    d = json.loads(l); k = d['roofline']['kernels_ms_per_step']
    print(round(d['value']), 'Mrays/s', round(d['ms_per_step'], 1), 'ms', {a: round(b, 1) for a, b in k.items()}, 'launches', d['gpu_launches'], 'paths', int(d['rays']['paths']), 'ext', int(d['rays']['ext']), 'shadow', int(d['rays']['shadow']), 'nodes/ray', round(d['roofline']['nodes_per_ray'], 2), 'tris/ray', round(d['roofline']['tris_per_ray'], 2), 'queued/ray', round(d['roofline']['bvh_rays_per_ray'], 3), 'bvh_ms', round(d['bvh_build_ms'], 1))
except Exception as e:
    print('FAILED', l[-300:])
")
  echo "$name $line" | tee -a $O/${TAG}_ab.txt
done < $SPEC
