#!/bin/bash
# One gpurun call: GPU tests, the default bench, the reference (CPU) arm, then -- plain run first -- an ncu launch list
# (time + DRAM bytes + lanes per launch of ONE profiled step -> profiles/rNN_traffic.json) and full captures of the two
# wavefront kernels.  Usage: tools/gpu_round.sh <tag> [skip-tests]
TAG=${1:-r02}
O=gpurun_out; mkdir -p $O
if [ "$2" != "skip-tests" ]; then
  python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest_gpu.log
fi
python bench.py > $O/${TAG}_bench_mesh1m.json 2> $O/${TAG}_bench_mesh1m.err; echo "bench rc=$?" >> $O/${TAG}_bench_mesh1m.err
python bench.py --impl reference --steps 4 --warmup 1 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct
T="python tools/ncu_target.py mesh1m 16 $O/${TAG}_ncu_rays_mesh1m.json"
$T > $O/${TAG}_target_plain.log 2>&1 &&
ncu --profile-from-start off --metrics $M --clock-control none --csv --log-file $O/${TAG}_launches_mesh1m.csv $T > $O/${TAG}_ncu_list.log 2>&1 &&
python tools/traffic_json.py $O/${TAG}_traffic.json mesh1m:$O/${TAG}_launches_mesh1m.csv:$O/${TAG}_ncu_rays_mesh1m.json > $O/${TAG}_traffic.log 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:^wf_traverse -s ${SKIP:-3} -c 2 -f -o $O/${TAG}_wf_traverse $T > $O/${TAG}_ncu_trav.log 2>&1 &&
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:^wf_logic -s ${SKIP:-3} -c 2 -f -o $O/${TAG}_wf_logic $T > $O/${TAG}_ncu_logic.log 2>&1
echo "ncu rc=$?" >> $O/${TAG}_ncu_list.log
