#!/bin/bash
# round 2, GPU call 21 (eight B200s): the final library under the driver's scaling protocol (N = 1, 2, 4, 8 at --steps 20
# --warmup 5, each with the single-process e2e leg), and the group tests (incl. block power) at 8 ranks
mkdir -p gpurun_out/r02
nproc > gpurun_out/r02/c21_box.txt
for n in 1 2 4 8; do
  if [ $n -eq 1 ]; then timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02/c21_bench_c5_n$n.json 2> gpurun_out/r02/c21_bench_n$n.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29800 + n)) bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r02/c21_bench_c5_n$n.json 2> gpurun_out/r02/c21_bench_n$n.err; fi
  python - gpurun_out/r02/c21_bench_c5_n$n.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], 'value', round(d['value'],1), 'us/trip', round(d['ms_per_step']*1e3,2), 'e2e', d.get('e2e') and (round(d['e2e']['value'],1), round(d['e2e']['seconds'],4)), 'parity', d.get('parity') and (d['parity']['lambda_rel_err'], d['parity']['ok']))
except Exception as e:
    print(sys.argv[1], 'FAILED', e)
PY
done
timeout 600 python -m pytest tests/test_local_group.py -m gpu -q -k "8" > gpurun_out/r02/c21_pytest_group8.log 2>&1; echo "rc=$?" >> gpurun_out/r02/c21_pytest_group8.log
tail -3 gpurun_out/r02/c21_pytest_group8.log
EIGSOL_B200_NGPU=8 EIGSOL_B200_TRACE=1 timeout 300 python tools/r02/e2e_trace.py > gpurun_out/r02/c21_e2e_trace_n8.log 2>&1; grep "== call" gpurun_out/r02/c21_e2e_trace_n8.log
