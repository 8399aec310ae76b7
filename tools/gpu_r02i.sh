#!/bin/bash
# round 2, final single-GPU call: the round script (tests, default bench, reference arm, ncu launch list -> traffic json, full captures of
# wf_traverse and wf_logic), then the pool-rule A/B and the other workloads at HEAD
bash tools/gpu_round.sh r02
O=gpurun_out
for k in wf_traverse wf_logic; do
  if [ -f $O/r02_$k.ncu-rep ]; then
    ncu -i $O/r02_$k.ncu-rep --page raw --csv > $O/r02_${k}_raw.csv 2>/dev/null
    ncu -i $O/r02_$k.ncu-rep --page details > $O/r02_${k}_details.txt 2>/dev/null
    ncu -i $O/r02_$k.ncu-rep --page source --csv > $O/r02_${k}_source.csv 2>/dev/null
  fi
done
bash tools/gpu_ab.sh r02i tools/ab/r02i.spec
