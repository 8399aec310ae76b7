#!/bin/bash
# round r01x: parity of the routed form of wf_logic, then routed vs unrouted on configs 3, 2, 5
O=gpurun_out
python -m pytest tests -m gpu -x -q -k "scheduling_forms or pool_and_chunk" > $O/r01x_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/r01x_ab.txt
tail -3 $O/r01x_pytest.log
tools/gpu_ab.sh r01x tools/ab_r01x.spec
