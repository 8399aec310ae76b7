#!/bin/bash
# round r01t: parity of the static work share, then the A/B list
O=gpurun_out
for st in 6 8; do
  CRT_WF_STATIC=$st python -m pytest tests -m gpu -x -q -k "pool_and_chunk or triangle_scene_radiance or frame_by_frame or virtual_ranks" > $O/r01t_pytest_static$st.log 2>&1; echo "static$st pytest rc=$?" | tee -a $O/r01t_ab.txt
done
for v in l2all tri2; do
  CRT_LIBCRT=$PWD/variants/libcrt_$v.so python -m pytest tests -m gpu -x -q -k "triangle_scene or soup_200k or golden" > $O/r01t_pytest_$v.log 2>&1; echo "$v pytest rc=$?" | tee -a $O/r01t_ab.txt
done
tools/gpu_ab.sh r01t tools/ab_r01t.spec
