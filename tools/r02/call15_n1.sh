#!/bin/bash
# round 2, GPU call 15 (one B200): the symmetric / Hermitian column-dot GEMV
mkdir -p gpurun_out/r02
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_block.py -m gpu -x -q -k "dense or Dense or block" > gpurun_out/r02/c15_pytest_dense.log 2>&1; echo "rc=$?" >> gpurun_out/r02/c15_pytest_dense.log
tail -8 gpurun_out/r02/c15_pytest_dense.log
one() { w=$1; shift; env "$@" timeout 300 python bench.py --workload $w --steps 200 --warmup 20 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w $*', round(d['ms_per_step']*1e3,2),'us', round(d['roofline']['frac'],4), d['launch'])"; }
{ one c3 EIGSOL_B200_GEMV_SYM=1; one c3 EIGSOL_B200_GEMV_SYM=0; one c3 EIGSOL_B200_GEMV_SYM=1; } 2>&1 | tee gpurun_out/r02/c15_gemv_sym.txt
timeout 300 python bench.py --workload c3 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02/c15_bench_c3.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/r02/c15_bench_c3.json').read().strip().splitlines()[-1]); print('c3 20/5', d['value'], d['ms_per_step'], d['roofline']['frac'], 'e2e', d['e2e']['value'], 'parity', d['parity'])"
