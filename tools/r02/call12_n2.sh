#!/bin/bash
# round 2, GPU call 12 (two B200s): group tests + trace after reverting the small allocations to plain memory; bench N=2
mkdir -p gpurun_out/r02
timeout 900 python -m pytest tests/test_local_group.py tests/test_dropin_cpp.py -m gpu -x -q > gpurun_out/r02/c12_pytest_group.log 2>&1; echo "rc=$?" >> gpurun_out/r02/c12_pytest_group.log
tail -4 gpurun_out/r02/c12_pytest_group.log
EIGSOL_B200_NGPU=2 EIGSOL_B200_TRACE=1 timeout 600 python tools/r02/e2e_trace.py > gpurun_out/r02/c12_e2e_trace_n2.log 2>&1
grep -E "== call" gpurun_out/r02/c12_e2e_trace_n2.log
awk '/== call 1/{f=1} f' gpurun_out/r02/c12_e2e_trace_n2.log | head -34
EIGSOL_B200_NGPU=2 timeout 300 python tools/r02/e2e_trace.py 2>&1 | grep "== call"
EIGSOL_B200_NGPU=1 timeout 300 python tools/r02/e2e_trace.py 2>&1 | grep "== call"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02/c12_bench_c5_n2_20.json 2> gpurun_out/r02/c12_bench_n2.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r02/c12_bench_c5_n2_20.json').read().strip().splitlines()[-1]); print('N=2', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['seconds'], 'parity', d['parity']['ok'])"
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N=1', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['seconds'])"
