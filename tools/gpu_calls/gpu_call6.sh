set -x
nvidia-smi -L
nvidia-smi topo -m | head -12
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/pytest_mgpu.log 2>&1; echo "rc=$?"; tail -30 gpurun_out/pytest_mgpu.log
for ex in peer nccl; do
EIGSOL_B200_EXCHANGE=$ex timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 200 --warmup 20 > gpurun_out/bench_c5_n2_$ex.json 2> gpurun_out/bench_c5_n2_$ex.err; echo "rc=$?"; cat gpurun_out/bench_c5_n2_$ex.json; tail -3 gpurun_out/bench_c5_n2_$ex.err
done
for w in c2 c3 c4; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --workload $w --no-e2e > gpurun_out/bench_${w}_n2.json 2> gpurun_out/bench_${w}_n2.err; echo "rc=$?"; cat gpurun_out/bench_${w}_n2.json; tail -3 gpurun_out/bench_${w}_n2.err
done
