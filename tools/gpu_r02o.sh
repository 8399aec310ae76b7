#!/bin/bash
# round 2 (gpurun --gpus 8; r02o = the same list with the two-lane device loop, plus N = 4 and N = 1 on the same box): strong scaling with the collective in every step on 8 B200s -- config 3 (split-spp and tiles), config 4
# (4K, 1024 spp per step), config 5 -- and the single-process group over 8 devices
O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/r02o_smi.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29621"
timeout 600 $TR bench.py --gpus 8 --steps 8 --warmup 3 > $O/r02o_bench_n8_strong_spp.json 2> $O/r02o_bench_n8_strong_spp.err
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29623"
timeout 600 $TR4 bench.py --gpus 4 --steps 8 --warmup 3 > $O/r02o_bench_n4_strong_spp.json 2> $O/r02o_bench_n4_strong_spp.err
timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-also > $O/r02o_bench_n1.json 2> $O/r02o_bench_n1.err
timeout 600 $TR bench.py --gpus 8 --steps 8 --warmup 3 --shard tiles > $O/r02o_bench_n8_strong_tiles.json 2> $O/r02o_bench_n8_strong_tiles.err
timeout 900 $TR bench.py --gpus 8 --steps 4 --warmup 3 --width 3840 --height 2160 --spp-per-step 1024 > $O/r02o_bench_n8_cfg4_4k_1024spp.json 2> $O/r02o_bench_n8_cfg4_4k_1024spp.err
