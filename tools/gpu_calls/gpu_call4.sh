set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu4.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/pytest_gpu4.log
for ex in peer nccl; do
EIGSOL_B200_EXCHANGE=$ex timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 200 --warmup 20 --no-e2e > gpurun_out/bench_c5_n2_$ex.json 2> gpurun_out/bench_c5_n2_$ex.err; echo "rc=$?"; cat gpurun_out/bench_c5_n2_$ex.json; tail -3 gpurun_out/bench_c5_n2_$ex.err
done
python bench.py --no-cpu-baseline --no-e2e > gpurun_out/bench_c5_n1_b.json 2>&1; cat gpurun_out/bench_c5_n1_b.json
