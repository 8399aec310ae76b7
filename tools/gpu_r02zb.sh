#!/bin/bash
# round 2: libcrt against the oracle on random scenes (tests/test_wgsl_literal.py::test_cuda_path_random_scenes) and the file's other GPU tests
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_wgsl_literal.py -m gpu -q > $O/r02zb_pytest_literal.log 2>&1; echo "rc=$?" >> $O/r02zb_pytest_literal.log
