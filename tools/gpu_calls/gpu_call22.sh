timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/pytest_gpu22.log 2>&1; echo "rc=$?"; tail -25 gpurun_out/pytest_gpu22.log
for w in c5 c2 c3 c4; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --workload $w > gpurun_out/bench22_${w}_n2.json 2> gpurun_out/bench22_${w}_n2.err; echo "rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench22_${w}_n2.json')); print('$w', d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e'] and d['e2e']['value'])"
done
