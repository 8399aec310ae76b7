#!/bin/bash
# A/B: traversal stack in shared memory, L1 prefetch of a group's later triangles
O=gpurun_out
CRT_LIBCRT=$PWD/variants/libcrt_ss8tpf.so python -m pytest tests -m gpu -x -q -k "triangle_scene or soup_200k or million_triangle or tie_rule" > $O/r02a_pytest.log 2>&1; echo "ss8tpf pytest rc=$?" | tee -a $O/r02a_ab.txt
tail -3 $O/r02a_pytest.log
tools/gpu_ab.sh r02a tools/ab_r02a.spec
