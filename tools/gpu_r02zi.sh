#!/bin/bash
# round 2, last call: smoke at HEAD, then a short default bench
O=gpurun_out; mkdir -p $O
timeout 70 python __graft_entry__.py smoke > $O/r02zi_smoke.log 2>&1; echo "smoke rc=$?" >> $O/r02zi_smoke.log
timeout 50 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-also > $O/r02zi_bench_short.json 2> $O/r02zi_bench_short.err
