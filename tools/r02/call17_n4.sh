#!/bin/bash
# round 2, GPU call 17 (four B200s): group + process-per-GPU parity (incl. the multi-GPU block power iteration) at world 4
# -- ranks with TWO neighbours -- and the driver-protocol bench at N=4
mkdir -p gpurun_out/r02
timeout 900 python -m pytest tests/test_local_group.py tests/test_dropin_cpp.py -m gpu -q -rs -k "4 or group or cache" > gpurun_out/r02/c17_pytest_group4.log 2>&1; echo "rc=$?" >> gpurun_out/r02/c17_pytest_group4.log
tail -6 gpurun_out/r02/c17_pytest_group4.log | cut -c1-200
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -q -k "4" > gpurun_out/r02/c17_pytest_mgpu4.log 2>&1; echo "rc=$?" >> gpurun_out/r02/c17_pytest_mgpu4.log
tail -4 gpurun_out/r02/c17_pytest_mgpu4.log | cut -c1-200
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29641 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/r02/c17_bench_c5_n4.json 2> gpurun_out/r02/c17_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r02/c17_bench_c5_n4.json').read().strip().splitlines()[-1]); print('N=4', d['value'], d['ms_per_step'], 'e2e', d['e2e'] and (d['e2e']['value'], d['e2e']['seconds']), 'parity', d['parity'] and d['parity']['ok'])"
EIGSOL_B200_NGPU=4 EIGSOL_B200_TRACE=1 timeout 300 python tools/r02/e2e_trace.py > gpurun_out/r02/c17_e2e_trace_n4.log 2>&1; grep "== call" gpurun_out/r02/c17_e2e_trace_n4.log
