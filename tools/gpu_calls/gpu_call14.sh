timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/pytest_gpu14.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/pytest_gpu14.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/latency_bench.py 16 64 128 256 2>/dev/null | grep '^{' | cut -c1-120
