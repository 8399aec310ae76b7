#!/bin/bash
# round 2, call 4 (gpurun --gpus 2): NCCL behind the C ABI and the single-process group on two physical GPUs; strong scaling with the
# collective inside every step
O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/r02d_smi.txt
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q --maxfail=8 -k "two_gpus or group_on_one" > $O/r02d_pytest_2gpu.log 2>&1; echo "rc=$?" >> $O/r02d_pytest_2gpu.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-also > $O/r02d_bench_n1.json 2> $O/r02d_bench_n1.err
NCCL_DEBUG=WARN timeout 600 $TR bench.py --gpus 2 --steps 4 --warmup 3 > $O/r02d_bench_n2_strong_spp.json 2> $O/r02d_bench_n2_strong_spp.err
timeout 600 $TR bench.py --gpus 2 --steps 4 --warmup 3 --shard tiles > $O/r02d_bench_n2_strong_tiles.json 2> $O/r02d_bench_n2_strong_tiles.err
timeout 600 $TR bench.py --gpus 2 --steps 4 --warmup 3 --scaling weak > $O/r02d_bench_n2_weak.json 2> $O/r02d_bench_n2_weak.err
timeout 600 $TR bench.py --gpus 2 --steps 4 --warmup 3 --spp-per-step 64 > $O/r02d_bench_n2_strong_spp64.json 2> $O/r02d_bench_n2_strong_spp64.err
timeout 300 $TR bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > $O/r02d_bench_n2_reference.json 2> $O/r02d_bench_n2_reference.err
