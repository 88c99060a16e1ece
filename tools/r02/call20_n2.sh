#!/bin/bash
# round 2, GPU call 20 (two B200s): bench N=2 after letting the other ranks leave before the e2e leg; C3 / C4 too
mkdir -p gpurun_out/r02
for w in c5 c4 c3; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 5 --workload $w > gpurun_out/r02/c20_bench_${w}_n2.json 2> gpurun_out/r02/c20_bench_${w}.err; echo "bench $w rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r02/c20_bench_${w}_n2.json').read().strip().splitlines()[-1]); print('$w N=2', d['value'], d['ms_per_step'], 'e2e', d['e2e'] and (d['e2e']['value'], d['e2e']['seconds']), 'parity', d['parity'] and d['parity']['ok'])"
tail -2 gpurun_out/r02/c20_bench_${w}.err | cut -c1-200
done
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('c5 N=1', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['seconds'])"
timeout 600 python -m pytest tests/test_local_group.py -m gpu -x -q -k "malformed or block" 2>&1 | tail -2
