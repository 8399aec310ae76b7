#!/bin/bash
# round r01w: parity of the routed form of wf_logic, then routed vs unrouted on configs 3, 2, 5
O=gpurun_out
python -m pytest tests -m gpu -x -q -k "scheduling_forms or pool_and_chunk" > $O/r01w_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/r01w_ab.txt
tail -3 $O/r01w_pytest.log
tools/gpu_ab.sh r01w tools/ab_r01w.spec
