python tools/latency_bench.py 16 64 128 > gpurun_out/lat_n1.jsonl 2>gpurun_out/lat_n1.err; cat gpurun_out/lat_n1.jsonl
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/latency_bench.py 16 64 128 256"
for dbg in 0 1 2 3 4 8 12 15; do
EIGSOL_B200_PEER_DEBUG=$dbg timeout 200 $T 2>gpurun_out/lat_n2_$dbg.err | grep '^{' | tee gpurun_out/lat_n2_$dbg.jsonl
done
EIGSOL_B200_EXCHANGE=nccl timeout 200 $T 2>gpurun_out/lat_n2_nccl.err | grep '^{' | tee gpurun_out/lat_n2_nccl.jsonl
