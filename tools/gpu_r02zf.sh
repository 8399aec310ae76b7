#!/bin/bash
# round 2, check at HEAD after the exact occluded-NEE contribution: smoke, the whole GPU suite in pytest -x order, the default bench, the reference arm
O=gpurun_out; mkdir -p $O
timeout 300 python __graft_entry__.py smoke > $O/r02zf_smoke.log 2>&1; echo "smoke rc=$?" >> $O/r02zf_smoke.log
timeout 2400 python -m pytest tests -m gpu -x -q > $O/r02zf_pytest_all.log 2>&1; echo "rc=$?" >> $O/r02zf_pytest_all.log
timeout 900 python bench.py > $O/r02zf_bench_default.json 2> $O/r02zf_bench_default.err; echo "rc=$?" >> $O/r02zf_bench_default.err
timeout 600 python bench.py --impl reference --steps 4 --warmup 1 > $O/r02zf_bench_reference.json 2> $O/r02zf_bench_reference.err
