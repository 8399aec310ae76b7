#!/bin/bash
# round 2, call 2: first GPU contact of the device-side loop, the group context, the device BVH build, the new tests
O=gpurun_out; mkdir -p $O
./variants/cond_graph_probe > $O/r02b_cond_probe.txt 2>&1
timeout 300 python __graft_entry__.py smoke > $O/r02b_smoke.log 2>&1; echo "smoke rc=$?" >> $O/r02b_smoke.log
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q --maxfail=8 -k "scheduling or dispatch_returns or builders or falls_back or far_from or adversarial or group_on_one" > $O/r02b_pytest_new.log 2>&1; echo "rc=$?" >> $O/r02b_pytest_new.log
timeout 1500 python -m pytest tests -m gpu -q --maxfail=8 --deselect tests/test_gpu_parity.py::test_wavefront_scheduling_forms_never_change_results > $O/r02b_pytest_all.log 2>&1; echo "rc=$?" >> $O/r02b_pytest_all.log
for b in host device; do
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-also --bvh $b > $O/r02b_bench_mesh1m_$b.json 2> $O/r02b_bench_mesh1m_$b.err
  timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-also --bvh $b --workload soup10m > $O/r02b_bench_soup10m_$b.json 2> $O/r02b_bench_soup10m_$b.err
done
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-also --wf-tuning 16 > $O/r02b_bench_mesh1m_hostloop.json 2> $O/r02b_bench_mesh1m_hostloop.err
timeout 900 python bench.py > $O/r02b_bench_default.json 2> $O/r02b_bench_default.err
