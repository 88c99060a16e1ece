nvidia-smi -L | head -2
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu8.log 2>&1; echo "rc=$?"; tail -25 gpurun_out/pytest_gpu8.log
python tools/latency_bench.py 16 64 128 256 2>&1 | grep '^{' | tee gpurun_out/lat8_n1.jsonl
EIGSOL_B200_SERPENTINE=0 python tools/latency_bench.py 128 256 2>&1 | grep '^{' | tee gpurun_out/lat8_n1_noserp.jsonl
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/latency_bench.py 16 64 128 256"
timeout 200 $T 2>gpurun_out/lat8_n2.err | grep '^{' | tee gpurun_out/lat8_n2.jsonl
EIGSOL_B200_EXCHANGE=nccl timeout 200 $T 2>gpurun_out/lat8_n2_nccl.err | grep '^{' | tee gpurun_out/lat8_n2_nccl.jsonl
tail -5 gpurun_out/lat8_n2.err
