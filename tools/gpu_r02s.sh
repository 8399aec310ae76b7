#!/bin/bash
# round 2: (1) the GPU tests of the second shader reading (tests/test_wgsl_literal.py); (2) wf_logic occupancy A/B: 4 CTAs of 256
# threads per SM (64 registers, 192 B spills), 7 or 6 CTAs of 128 threads (72 / 80 registers) against 3 x 256 (80 registers)
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_wgsl_literal.py -m gpu -x -q > $O/r02s_pytest_literal.log 2>&1; echo "rc=$?" >> $O/r02s_pytest_literal.log
timeout 1500 bash tools/gpu_ab.sh r02s tools/ab/r02s.spec
