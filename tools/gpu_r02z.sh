#!/bin/bash
# round 2 (gpurun --gpus 2), HEAD after the record prefetch and the counter blocks: the 2-GPU tests (NCCL behind the C ABI, single-process
# group) that a 1-GPU box skips, then N = 1 and N = 2 strong scaling with the collective inside every step on the same box
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -q --maxfail=8 -k "two_gpus or group_on_one or multi" > $O/r02z_pytest_2gpu.log 2>&1; echo "rc=$?" >> $O/r02z_pytest_2gpu.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-also > $O/r02z_bench_n1.json 2> $O/r02z_bench_n1.err
timeout 300 $TR bench.py --gpus 2 --steps 4 --warmup 3 > $O/r02z_bench_n2_strong_spp.json 2> $O/r02z_bench_n2_strong_spp.err
