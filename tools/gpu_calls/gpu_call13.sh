export LAT_PER_RANK=1
export LAT_CONFIGS="default:;nowait:EIGSOL_B200_DEBUG_NOWAIT=1;nosend:EIGSOL_B200_DEBUG_NOSEND=1;nowait_nosend:EIGSOL_B200_DEBUG_NOWAIT=1,EIGSOL_B200_DEBUG_NOSEND=1;noserp:EIGSOL_B200_SERPENTINE=0"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 tools/latency_bench.py 256 128 2>gpurun_out/lat13.err | grep -E '^\{|^#' | cut -c1-130 | tee gpurun_out/lat13_n8.txt
