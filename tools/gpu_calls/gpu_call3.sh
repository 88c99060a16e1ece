set -x
nvidia-smi -L
python -m pytest tests/test_multi_gpu.py tests/test_dropin_cpp.py -m gpu -x -q > gpurun_out/pytest_mgpu.log 2>&1; echo "rc=$?"; tail -30 gpurun_out/pytest_mgpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 200 --warmup 20 > gpurun_out/bench_c5_n2.json 2> gpurun_out/bench_c5_n2.err; echo "rc=$?"; cat gpurun_out/bench_c5_n2.json; tail -5 gpurun_out/bench_c5_n2.err
python tools/debug_sell.py > gpurun_out/debug_sell.log 2>&1; cat gpurun_out/debug_sell.log
