#!/bin/bash
# round 2, GPU call 1 (one B200): full GPU test suite, bench C5 at the driver's protocol, setup trace of the
# one-shot call, variant sweep on the non-stencil shapes, read-mix HBM probe.
mkdir -p gpurun_out/r02
nvidia-smi -L > gpurun_out/r02/c01_gpus.txt
nproc >> gpurun_out/r02/c01_gpus.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02/c01_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02/c01_pytest.log
tail -5 gpurun_out/r02/c01_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02/c01_bench_c5.json 2> gpurun_out/r02/c01_bench_c5.err; echo "bench rc=$?"
cat gpurun_out/r02/c01_bench_c5.json
timeout 600 python bench.py --steps 200 --warmup 20 --no-cpu-baseline > gpurun_out/r02/c01_bench_c5_200.json 2>> gpurun_out/r02/c01_bench_c5.err
cat gpurun_out/r02/c01_bench_c5_200.json
EIGSOL_B200_TRACE=1 timeout 600 python tools/r02/e2e_trace.py > gpurun_out/r02/c01_e2e_trace.log 2>&1
tail -40 gpurun_out/r02/c01_e2e_trace.log
python -c "
import pcsc_eigenvalue_solver_project_b200 as es
for planes in (1, 4, 8, 12):
    print('hbm_probe planes', planes, es.hbm_probe(128 << 20, planes, 30))
" > gpurun_out/r02/c01_hbm_probe.txt 2>&1
cat gpurun_out/r02/c01_hbm_probe.txt
TRIPS=100 timeout 900 python tools/variant_bench.py s16 s32 s64 i8 i32 > gpurun_out/r02/c01_variants_f64.jsonl 2> gpurun_out/r02/c01_variants.err
tail -3 gpurun_out/r02/c01_variants_f64.jsonl
for w in c2 c3 c4; do timeout 300 python bench.py --workload $w --steps 200 --warmup 20 --no-cpu-baseline > gpurun_out/r02/c01_bench_$w.json 2>> gpurun_out/r02/c01_bench_c5.err; cat gpurun_out/r02/c01_bench_$w.json; done
