#!/bin/bash
# One gpurun call at HEAD: GPU tests, default bench, reference (CPU) arm, config 2 (Cornell, AUTO -> wavefront), then --
# plain run first -- the ncu launch list of a short Cornell bench.  Usage: tools/gpu_round_s.sh <tag>
TAG=${1:-r01s}
O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest_gpu.log
python bench.py > $O/${TAG}_bench_mesh1m.json 2> $O/${TAG}_bench_mesh1m.err; echo "bench rc=$?" >> $O/${TAG}_bench_mesh1m.err
python bench.py --impl reference --steps 4 --warmup 1 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err
python bench.py --workload cornell --steps 8 --warmup 3 --no-also > $O/${TAG}_cfg2_cornell_1080p.json 2> $O/${TAG}_cfg2.err
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct
SHORT="python bench.py --workload cornell --steps 1 --warmup 1 --spp-per-step 16 --no-cpu-baseline --no-also"
$SHORT > $O/${TAG}_short_plain.log 2>&1 &&
ncu --metrics $M --clock-control none -c 400 --csv --log-file $O/${TAG}_launches_cornell.csv $SHORT > $O/${TAG}_ncu_list.log 2>&1
echo "ncu rc=$?" >> $O/${TAG}_ncu_list.log
tail -3 $O/${TAG}_pytest_gpu.log; cut -c1-300 $O/${TAG}_bench_mesh1m.json
