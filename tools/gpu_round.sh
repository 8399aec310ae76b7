#!/bin/bash
# One gpurun call: GPU tests, the default bench, the reference (CPU) arm, then -- plain run first -- an ncu launch list
# and full captures of the two wavefront kernels on a short bench.  Usage: tools/gpu_round.sh <tag>
TAG=${1:-r01}
O=gpurun_out
SHORT="python bench.py --steps 1 --warmup 1 --spp-per-step 16 --no-cpu-baseline --no-also"
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest_gpu.log
python bench.py > $O/${TAG}_bench_mesh1m.json 2> $O/${TAG}_bench_mesh1m.err; echo "bench rc=$?" >> $O/${TAG}_bench_mesh1m.err
python bench.py --impl reference --steps 4 --warmup 1 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct
$SHORT > $O/${TAG}_short_plain.log 2>&1 &&
ncu --metrics $M --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv $SHORT > $O/${TAG}_ncu_list.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:^wf_traverse$ -s ${SKIP:-33} -c 2 -f -o $O/${TAG}_wf_traverse $SHORT > $O/${TAG}_ncu_trav.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:^wf_logic$ -s ${SKIP:-33} -c 2 -f -o $O/${TAG}_wf_logic $SHORT > $O/${TAG}_ncu_logic.log 2>&1
echo "ncu rc=$?" >> $O/${TAG}_ncu_list.log
