#!/bin/bash
# round 2: wf_logic's next-trip record prefetch as the default for scenes with a BVH (template parameter PF): parity subset, then the
# bench workloads at HEAD (compare with base / spp32_base / cornell_base / soup_base of r02u / r02r)
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_wgsl_literal.py -m gpu -x -q -k "scheduling or triangle_lights or full_size or dispatch_returns or error_behaviour or literal" > $O/r02v_pytest_subset.log 2>&1; echo "rc=$?" >> $O/r02v_pytest_subset.log
timeout 900 bash tools/gpu_ab.sh r02v tools/ab/r02v.spec
