#!/bin/bash
# round 2, call 7: triangle lights in the wavefront form (tests), pool sweep for short chunks, coherence bound, Cornell single frames,
# compute-sanitizer
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q --maxfail=8 -k "triangle_lights" > $O/r02g_pytest_trilights.log 2>&1; echo "rc=$?" >> $O/r02g_pytest_trilights.log
timeout 1800 python -m pytest tests -m gpu -q --maxfail=8 > $O/r02g_pytest_all.log 2>&1; echo "rc=$?" >> $O/r02g_pytest_all.log
bash tools/gpu_ab.sh r02g tools/ab/r02g.spec
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --workload cornell --kernel wavefront > $O/r02g_bench_cornell_wavefront.json 2> $O/r02g_bench_cornell_wavefront.err
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --workload cornell > $O/r02g_bench_cornell_auto.json 2> $O/r02g_bench_cornell_auto.err
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python tools/sanitize_target.py > $O/r02g_sanitizer_memcheck.log 2>&1; echo "memcheck rc=$?" >> $O/r02g_sanitizer_memcheck.log
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python tools/sanitize_target.py > $O/r02g_sanitizer_racecheck.log 2>&1; echo "racecheck rc=$?" >> $O/r02g_sanitizer_racecheck.log
