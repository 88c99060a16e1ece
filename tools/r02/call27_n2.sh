#!/bin/bash
# round 2, GPU call 27 (two B200s): bench with the e2e leg of N > 1 made from a clean child process
mkdir -p gpurun_out/r02
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02/c27_bench_c5_n2.json 2> gpurun_out/r02/c27_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r02/c27_bench_c5_n2.json').read().strip().splitlines()[-1]); print('N=2', d['value'], d['ms_per_step'], 'e2e', d['e2e'] and (d['e2e']['value'], d['e2e']['seconds']), 'parity', d['parity'])"
grep "bench.py:" gpurun_out/r02/c27_bench.err | tail -3
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/r02/c27_bench_c5_n1.json 2>> gpurun_out/r02/c27_bench.err; python -c "
import json; d=json.loads(open('gpurun_out/r02/c27_bench_c5_n1.json').read().strip().splitlines()[-1]); print('N=1', d['value'], d['ms_per_step'], 'e2e', d['e2e'] and (d['e2e']['value'], d['e2e']['seconds']), 'parity', d['parity'] and d['parity']['ok'], 'cpu', d['cpu_baseline']['value'])"
timeout 300 python bench.py --workload c3 --steps 20 --warmup 5 --no-cpu-baseline | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('c3 N=1', d['value'], 'e2e', d['e2e'] and d['e2e']['value'], d['parity'] and d['parity']['ok'])"
