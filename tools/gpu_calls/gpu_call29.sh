nvidia-smi -L | wc -l
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q -k "8-peer or 4-nccl" > gpurun_out/pytest_mgpu29.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/pytest_mgpu29.log | cut -c1-300
for w in c5 c4 c3; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --workload $w $( [ $w != c5 ] && echo --no-e2e ) > gpurun_out/bench29_${w}_n8.json 2> gpurun_out/bench29_${w}_n8.err; echo "rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench29_${w}_n8.json')); print('$w', d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e'] and d['e2e']['value'], d['clocks'])"
done
LAT_PER_RANK=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 tools/latency_bench.py 256 2>/dev/null | grep -E '^\{|^#' | cut -c1-130
