timeout 900 python -m pytest tests/test_multi_gpu.py tests/test_gpu_parity.py -m gpu -x -q -k "row_sharded or dense" > gpurun_out/pytest_gpu24.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/pytest_gpu24.log
for w in c3; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --workload $w --no-e2e 2>/dev/null > gpurun_out/bench24_${w}_n2.json; python -c "
import json; d=json.load(open('gpurun_out/bench24_${w}_n2.json')); print('$w', d['value'], d['ms_per_step'], d['roofline']['frac'])"
done
