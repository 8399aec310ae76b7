#!/bin/bash
# round 2 (gpurun --gpus 8): the N = 8 lines again with the timing loop that no longer waits on the host between steps
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29631"
timeout 600 $TR bench.py --gpus 8 --steps 8 --warmup 3 > $O/r02l_bench_n8_strong_spp.json 2> $O/r02l_bench_n8_strong_spp.err
timeout 600 $TR bench.py --gpus 8 --steps 8 --warmup 3 --shard tiles > $O/r02l_bench_n8_strong_tiles.json 2> $O/r02l_bench_n8_strong_tiles.err
TR2="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29633"
timeout 600 $TR2 bench.py --gpus 2 --steps 8 --warmup 3 > $O/r02l_bench_n2_strong_spp.json 2> $O/r02l_bench_n2_strong_spp.err
