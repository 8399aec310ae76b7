#!/bin/bash
# ncu --set full captures of the wavefront kernels on a short bench (plain run first).  Usage: tools/gpu_prof.sh <tag> [kernels...]
TAG=${1:-prof}; shift
KERNELS=${@:-wf_shade wf_regen wf_traverse}
O=gpurun_out
SHORT="python bench.py --steps 1 --warmup 1 --spp-per-step 16 --no-cpu-baseline --no-also"
$SHORT > $O/${TAG}_short_plain.log 2>&1 || { echo "plain run failed"; tail -5 $O/${TAG}_short_plain.log; exit 1; }
for K in $KERNELS; do
  ncu --set full --clock-control none --import-source on -k regex:^${K}$ -s ${SKIP:-66} -c 2 -f -o $O/${TAG}_$K $SHORT > $O/${TAG}_ncu_$K.log 2>&1
  echo "$K rc=$?"
done
