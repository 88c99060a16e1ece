#!/bin/bash
# round 2, GPU call 14 (one B200): full GPU suite, GEMV out-of-line combine, smoke, driver-protocol bench of every config
mkdir -p gpurun_out/r02
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02/c14_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02/c14_pytest.log
tail -5 gpurun_out/r02/c14_pytest.log
one() { w=$1; shift; env "$@" timeout 300 python bench.py --workload $w --steps 200 --warmup 20 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w $*', round(d['ms_per_step']*1e3,2),'us', round(d['roofline']['frac'],4))"; }
{ one c3 EIGSOL_B200_GEMV_VAR=0; one c3 EIGSOL_B200_GEMV_VAR=3; one c3 EIGSOL_B200_GEMV_VAR=3 EIGSOL_B200_GEMV_WAVES=4; one c3 EIGSOL_B200_GEMV_VAR=0; one c3 EIGSOL_B200_GEMV_VAR=3; } 2>&1 | tee gpurun_out/r02/c14_gemv.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02/c14_bench_c5.json 2> gpurun_out/r02/c14_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r02/c14_bench_ref.json 2>> gpurun_out/r02/c14_bench.err; echo "ref rc=$?"
cat gpurun_out/r02/c14_bench_c5.json gpurun_out/r02/c14_bench_ref.json | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print(d.get('impl','b200'), d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'cfg', d['config'], 'steps', d['steps'], d['warmup'])"
