#!/bin/bash
# round 2, GPU call 24 (two B200s): phase trace of the e2e leg INSIDE bench.py under torchrun (why 64 ms there, 44 ms alone?)
mkdir -p gpurun_out/r02
EIGSOL_B200_TRACE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02/c24_bench_c5_n2.json 2> gpurun_out/r02/c24_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r02/c24_bench_c5_n2.json').read().strip().splitlines()[-1]); print('N=2', d['value'], d['ms_per_step'], 'e2e', d['e2e'] and (d['e2e']['value'], d['e2e']['seconds']))"
grep "eigsol_b200\]" gpurun_out/r02/c24_bench.err | tail -40
