#!/bin/bash
# lazy L (L in its own sector, touched only by passes that add to it or finish the path): full GPU tests, then A/B against HEAD
O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/r02d_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/r02d_ab.txt
tail -3 $O/r02d_pytest_gpu.log
tools/gpu_ab.sh r02d tools/ab_r02d.spec
