#!/bin/bash
# round 2: no-triangle instantiation of wf_logic without any pending-NEE handling (0 spills): tests on linear-list scenes, Cornell bench
O=gpurun_out; mkdir -p $O
timeout 120 python -m pytest tests -m gpu -q -x -k "cornell or golden or literal or single_frame or linear" > $O/r02zh_pytest_subset.log 2>&1; echo "rc=$?" >> $O/r02zh_pytest_subset.log
cat > /tmp/zh.spec <<'EOS'
head_cornell | - | - | --workload cornell
EOS
timeout 100 bash tools/gpu_ab.sh r02zh /tmp/zh.spec
