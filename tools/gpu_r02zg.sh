#!/bin/bash
# round 2: no-BVH instantiation of wf_logic without the pending-NEE code (wf_bounce<.., BVH = false>): tests on linear-list scenes, Cornell and default bench
O=gpurun_out; mkdir -p $O
timeout 200 python -m pytest tests -m gpu -q -x -k "cornell or golden or literal or single_frame or linear" > $O/r02zg_pytest_subset.log 2>&1; echo "rc=$?" >> $O/r02zg_pytest_subset.log
cat > /tmp/zg.spec <<'EOS'
head_cornell | - | - | --workload cornell
head | - | - |
EOS
timeout 200 bash tools/gpu_ab.sh r02zg /tmp/zg.spec
