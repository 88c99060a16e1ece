#!/bin/bash
# round 2, GPU call 30 (two B200s): final check of bench.py at N=2 (in-process e2e leg, two warm-up calls)
mkdir -p gpurun_out/r02
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02/c30_bench_c5_n2.json 2> gpurun_out/r02/c30_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r02/c30_bench_c5_n2.json').read().strip().splitlines()[-1]); print('N=2', d['value'], d['ms_per_step'], 'e2e', d['e2e'] and (d['e2e']['value'], d['e2e']['seconds']), 'parity', d['parity'] and d['parity']['ok'], 'launches', d['gpu_launches'])"
grep "bench.py:" gpurun_out/r02/c30_bench.err | tail -3
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 2>/dev/null | cut -c1-200
