timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 > gpurun_out/bench31_c5_n8.json 2> gpurun_out/bench31_c5_n8.err; echo "rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench31_c5_n8.json')); print('c5', d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e'] and d['e2e']['value'], d['clocks'])"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 8 --steps 1000 --warmup 50 --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('c5 1000 steps', d['value'], d['ms_per_step'])"
