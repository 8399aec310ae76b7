#!/bin/bash
# round 2: after the tree-cache fix -- the subset that exposed it (in that order), the whole suite, the no-alloc file alone, soup A/B
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "scheduling or triangle_lights or full_size or dispatch_returns or error_behaviour or reused_address" > $O/r02n_pytest_subset.log 2>&1; echo "rc=$?" >> $O/r02n_pytest_subset.log
timeout 2400 python -m pytest tests -m gpu -x -q > $O/r02n_pytest_all.log 2>&1; echo "rc=$?" >> $O/r02n_pytest_all.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/r02n_pytest_noalloc_alone.log 2>&1; echo "rc=$?" >> $O/r02n_pytest_noalloc_alone.log
timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-also --workload soup10m > $O/r02n_bench_soup10m.json 2> $O/r02n_bench_soup10m.err
timeout 900 python bench.py > $O/r02n_bench_default.json 2> $O/r02n_bench_default.err
