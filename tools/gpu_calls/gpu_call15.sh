export LAT_CONFIGS="default:;nowait:EIGSOL_B200_DEBUG_NOWAIT=1"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 tools/latency_bench.py 16 128 256 2>gpurun_out/lat15.err | grep -E '^\{' | cut -c1-130 | tee gpurun_out/lat15_n8.txt
