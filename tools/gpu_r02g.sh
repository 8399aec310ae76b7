#!/bin/bash
# round 2, call 7: triangle lights in the wavefront form (tests), pool sweep for short chunks, coherence bound, Cornell single frames,
# compute-sanitizer
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q --maxfail=8 -k "triangle_lights" > $O/r02g_pytest_trilights.log 2>&1; echo "rc=$?" >> $O/r02g_pytest_trilights.log
timeout 1800 python -m pytest tests -m gpu -q --maxfail=8 > $O/r02g_pytest_all.log 2>&1; echo "rc=$?" >> $O/r02g_pytest_all.log
bash tools/gpu_ab.sh r02g tools/ab/r02g.spec
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --workload cornell --kernel wavefront > $O/r02g_bench_cornell_wavefront.json 2> $O/r02g_bench_cornell_wavefront.err
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --workload cornell > $O/r02g_bench_cornell_auto.json 2> $O/r02g_bench_cornell_auto.err
# compute-sanitizer is closed on this pool ("closed on this pool and stays closed": first r02g run), so its target runs plainly:
# every kernel form on a 96x64 mesh render, each compared bit for bit with the first
timeout 600 python tools/sanitize_target.py > $O/r02g_sanitize_target_plain.log 2>&1; echo "rc=$?" >> $O/r02g_sanitize_target_plain.log
