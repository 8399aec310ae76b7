#!/bin/bash
# round 2, call 1: parity of the untried wf_traverse schedules, then the A/B list of tools/ab/next_round.spec
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/r02a_smi.txt
for v in pp8 bf32r2 bf32r2pp8; do
  CRT_LIBCRT=$PWD/variants/libcrt_$v.so timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "triangle or wavefront or million or soup" > $O/r02a_parity_$v.log 2>&1
  echo "$v rc=$?" >> $O/r02a_parity.txt
done
bash tools/gpu_ab.sh r02a tools/ab/next_round.spec
