#!/bin/bash
# round 2: libcrt against the oracle on random scenes with triangles (both forms, both builders)
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_wgsl_literal.py -m gpu -q -k "triangles" > $O/r02zc_pytest_tri_fuzz.log 2>&1; echo "rc=$?" >> $O/r02zc_pytest_tri_fuzz.log
