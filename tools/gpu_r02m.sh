#!/bin/bash
# round 2: two half pools on two streams (the device loop's default) against one pool: tests first, then the A/B
O=gpurun_out; mkdir -p $O
timeout 300 python __graft_entry__.py smoke > $O/r02m_smoke.log 2>&1; echo "smoke rc=$?" >> $O/r02m_smoke.log
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "scheduling or triangle_lights or full_size or dispatch_returns or error_behaviour" > $O/r02m_pytest_lanes.log 2>&1; echo "rc=$?" >> $O/r02m_pytest_lanes.log
bash tools/gpu_ab.sh r02m tools/ab/r02m.spec
timeout 2400 python -m pytest tests -m gpu -x -q > $O/r02m_pytest_all.log 2>&1; echo "rc=$?" >> $O/r02m_pytest_all.log
timeout 600 python tools/sanitize_target.py > $O/r02m_sanitize_target_plain.log 2>&1; echo "rc=$?" >> $O/r02m_sanitize_target_plain.log
