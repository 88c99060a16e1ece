#!/bin/bash
# round 2, GPU call 6 (one B200): full GPU suite after the kernel / pool / cache changes, driver-protocol bench, GEMV and
# SpMV prefetch sweeps, e2e trace, finer dispatch sweep
mkdir -p gpurun_out/r02
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02/c06_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02/c06_pytest.log
tail -6 gpurun_out/r02/c06_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02/c06_bench_c5.json 2> gpurun_out/r02/c06_bench_c5.err; echo "bench rc=$?"; cat gpurun_out/r02/c06_bench_c5.json
one() {  # workload env...
  w=$1; shift
  env "$@" timeout 300 python bench.py --workload $w --steps 200 --warmup 20 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w $*', round(d['ms_per_step']*1e3,2),'us', round(d['roofline']['frac'],4))"
}
{
for pf in 0 8 16 32 64; do for wv in 4 5; do one c3 EIGSOL_B200_GEMV_PREFETCH=$pf EIGSOL_B200_GEMV_WAVES=$wv; done; done
for pf in 0 1 2; do one c5 EIGSOL_B200_SPMV_PREFETCH=$pf; one c2 EIGSOL_B200_SPMV_PREFETCH=$pf; BENCH_GRID=128 one c5 EIGSOL_B200_SPMV_PREFETCH=$pf; done
} 2>&1 | tee gpurun_out/r02/c06_prefetch.txt
EIGSOL_B200_TRACE=1 timeout 600 python tools/r02/e2e_trace.py > gpurun_out/r02/c06_e2e_trace.log 2>&1
tail -26 gpurun_out/r02/c06_e2e_trace.log
for w in c2 c3 c4; do timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02/c06_bench_$w.json 2>> gpurun_out/r02/c06_bench_c5.err; cat gpurun_out/r02/c06_bench_$w.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$w', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['parity'])"; done
TRIPS=100 timeout 900 python tools/variant_bench.py s4 s8 s12 i4 i12 i16 s16c i8c i32c > gpurun_out/r02/c06_variants.jsonl 2> gpurun_out/r02/c06_variants.err
tail -2 gpurun_out/r02/c06_variants.jsonl
