#!/bin/bash
# round 2, check at HEAD after the wf_logic record prefetch: smoke, the round script (whole GPU suite, default bench, reference arm, ncu
# launch list -> traffic json, full captures of wf_traverse and wf_logic) and the text pages of the captures
O=gpurun_out; mkdir -p $O
timeout 300 python __graft_entry__.py smoke > $O/r02w_smoke.log 2>&1; echo "smoke rc=$?" >> $O/r02w_smoke.log
bash tools/gpu_round.sh r02w
for k in wf_traverse wf_logic; do
  if [ -f $O/r02w_$k.ncu-rep ]; then
    ncu -i $O/r02w_$k.ncu-rep --page raw --csv > $O/r02w_${k}_raw.csv 2>/dev/null
    ncu -i $O/r02w_$k.ncu-rep --page source --csv > $O/r02w_${k}_source.csv 2>/dev/null
  fi
done
