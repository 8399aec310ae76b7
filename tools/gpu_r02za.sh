#!/bin/bash
# round 2: the GPU tests of tests/test_wgsl_literal.py (default scene and the two-light / glass-pane variant, both kernel forms)
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_wgsl_literal.py -m gpu -q > $O/r02za_pytest_literal.log 2>&1; echo "rc=$?" >> $O/r02za_pytest_literal.log
