#!/bin/bash
# round 2, call 3: wf_tail, PLOC radius 1, division-free uv test: tests, then the effect on short steps and single frames
O=gpurun_out; mkdir -p $O
timeout 300 python __graft_entry__.py smoke > $O/r02c_smoke.log 2>&1; echo "smoke rc=$?" >> $O/r02c_smoke.log
timeout 1800 python -m pytest tests -m gpu -q --maxfail=8 > $O/r02c_pytest_all.log 2>&1; echo "rc=$?" >> $O/r02c_pytest_all.log
for spp in 32 256; do
  for tail in 0 0x1f00 0xf00 0x1100 0x1200; do
    timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-also --spp-per-step $spp --wf-tuning $tail > $O/r02c_bench_spp${spp}_tail${tail}.json 2> $O/r02c_bench_spp${spp}_tail${tail}.err
  done
done
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-also --workload soup10m > $O/r02c_bench_soup10m.json 2> $O/r02c_bench_soup10m.err
timeout 900 python bench.py > $O/r02c_bench_default.json 2> $O/r02c_bench_default.err
