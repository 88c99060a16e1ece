#!/bin/bash
# round 2, GPU call 32 (eight B200s): bench line of the final tree at N=8, e2e leg included
mkdir -p gpurun_out/r02
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r02/c32_bench_c5_n8.json 2> gpurun_out/r02/c32_bench.err; echo "bench rc=$?"
