#!/bin/bash
# round 2: the allocation-free entry point -- its test alone in a fresh process, then after the whole GPU suite (pytest -x order)
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/r02h_pytest_noalloc_alone.log 2>&1; echo "rc=$?" >> $O/r02h_pytest_noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -x -q > $O/r02h_pytest_all.log 2>&1; echo "rc=$?" >> $O/r02h_pytest_all.log
timeout 600 python tools/sanitize_target.py > $O/r02h_sanitize_target_plain.log 2>&1; echo "rc=$?" >> $O/r02h_sanitize_target_plain.log
