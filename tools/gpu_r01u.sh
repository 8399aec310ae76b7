#!/bin/bash
# round r01u: parity of the routed form of wf_logic, then routed vs unrouted on configs 3, 2, 5
O=gpurun_out
python -m pytest tests -m gpu -x -q -k "scheduling_forms or pool_and_chunk" > $O/r01u_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/r01u_ab.txt
tail -3 $O/r01u_pytest.log
tools/gpu_ab.sh r01u tools/ab_r01u.spec
