#!/bin/bash
# round 2, GPU call 8 (one B200): full GPU suite, GEMV loop / mapping variants, dispatch sweep
mkdir -p gpurun_out/r02
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02/c08_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02/c08_pytest.log
tail -6 gpurun_out/r02/c08_pytest.log
one() {  # workload env...
  w=$1; shift
  env "$@" timeout 300 python bench.py --workload $w --steps 200 --warmup 20 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w $*', round(d['ms_per_step']*1e3,2),'us', round(d['roofline']['frac'],4))"
}
{
for var in 0 1; do for map in 0 1; do for wv in 4 5 6 8; do one c3 EIGSOL_B200_GEMV_VAR=$var EIGSOL_B200_GEMV_MAP=$map EIGSOL_B200_GEMV_WAVES=$wv; done; done; done
one c5 X=1; one c2 X=1; one c4 X=1
} 2>&1 | tee gpurun_out/r02/c08_gemv.txt
TRIPS=100 timeout 900 python tools/variant_bench.py s4 s8 s12 i4 i8 i12 i16 s16c i8c i32c > gpurun_out/r02/c08_variants.jsonl 2> gpurun_out/r02/c08_variants.err
tail -2 gpurun_out/r02/c08_variants.jsonl; tail -2 gpurun_out/r02/c08_variants.err
