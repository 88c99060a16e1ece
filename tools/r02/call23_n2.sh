#!/bin/bash
# round 2, GPU call 23 (two B200s): small buffers from the pool on the handle's stream -- full parity at 1 and 2 GPUs, traces
mkdir -p gpurun_out/r02
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r02/c23_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r02/c23_pytest.log
tail -3 gpurun_out/r02/c23_pytest.log
EIGSOL_B200_NGPU=2 EIGSOL_B200_TRACE=1 timeout 600 python tools/r02/e2e_trace.py > gpurun_out/r02/c23_e2e_trace_n2.log 2>&1
grep -E "== call" gpurun_out/r02/c23_e2e_trace_n2.log
awk '/== call 1/{f=1} f' gpurun_out/r02/c23_e2e_trace_n2.log | grep -E "streams|peer map|unit map|destroy|initialised|end \(" | head -12
EIGSOL_B200_NGPU=1 EIGSOL_B200_TRACE=1 timeout 300 python tools/r02/e2e_trace.py 2>&1 | grep -E "== call|create: handle|destroy" | tail -6
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02/c23_bench_c5_n2.json 2> gpurun_out/r02/c23_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r02/c23_bench_c5_n2.json').read().strip().splitlines()[-1]); print('N=2', d['value'], d['ms_per_step'], 'e2e', d['e2e'] and (d['e2e']['value'], d['e2e']['seconds']), 'parity', d['parity'] and d['parity']['ok'])"
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
