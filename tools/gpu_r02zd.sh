#!/bin/bash
# round 2: after the exact occluded-NEE contribution (nee_encode / nee_resolve): fuzz diagnostic, the literal file's GPU tests, perf check
O=gpurun_out; mkdir -p $O
timeout 300 python tools/fuzz_diag.py 8 > $O/r02zd_fuzz_diag.txt 2>&1
timeout 600 python -m pytest tests/test_wgsl_literal.py -m gpu -q > $O/r02zd_pytest_literal.log 2>&1; echo "rc=$?" >> $O/r02zd_pytest_literal.log
cat > /tmp/zd.spec <<'EOS'
head | - | - |
head_cornell | - | - | --workload cornell
EOS
timeout 600 bash tools/gpu_ab.sh r02zd /tmp/zd.spec
