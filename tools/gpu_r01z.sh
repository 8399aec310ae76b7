#!/bin/bash
# round r01z: the split form (wf_shade + wf_regen) over class-sorted lists against the fused wf_logic
O=gpurun_out
python -m pytest tests -m gpu -x -q -k "kernel_forms_agree or scheduling_forms" > $O/r01z_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/r01z_ab.txt
tail -3 $O/r01z_pytest.log
tools/gpu_ab.sh r01z tools/ab_r01z.spec
