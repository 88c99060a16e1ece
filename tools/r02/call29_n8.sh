#!/bin/bash
# round 2, GPU call 29 (eight B200s): N=1 and N=8 under the driver's protocol with whole-batch graphs (device-resident only)
mkdir -p gpurun_out/r02
for rep in 1 2; do
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline --no-e2e 2>/dev/null > gpurun_out/r02/c29_n1_$rep.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29808 bench.py --gpus 8 --steps 20 --warmup 5 --no-e2e 2>/dev/null > gpurun_out/r02/c29_n8_$rep.json
python - <<PY
import json
a=json.loads(open('gpurun_out/r02/c29_n1_$rep.json').read().strip().splitlines()[-1]); b=json.loads(open('gpurun_out/r02/c29_n8_$rep.json').read().strip().splitlines()[-1])
print('rep $rep N=1', round(a['value'],1), round(a['ms_per_step']*1e3,2), 'N=8', round(b['value'],1), round(b['ms_per_step']*1e3,2), 'ratio', round(b['value']/a['value'],3))
PY
done
