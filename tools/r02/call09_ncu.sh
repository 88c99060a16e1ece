#!/bin/bash
# round 2 (one B200): ncu launch list of the driver-protocol bench command and one full capture of the dominant kernel
mkdir -p gpurun_out/r02
CMD="python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/r02/c09_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02/c09_launches_c5.csv $CMD > gpurun_out/r02/c09_ncu_launches.log 2>&1
tail -3 gpurun_out/r02/c09_ncu_launches.log
$CMD > gpurun_out/r02/c09_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:spmv_csr_scalar -s 12 -c 2 -o gpurun_out/r02/c09_full_c5 $CMD > gpurun_out/r02/c09_ncu_full.log 2>&1
tail -3 gpurun_out/r02/c09_ncu_full.log
CMD3="python bench.py --workload c3 --steps 20 --warmup 5 --no-cpu-baseline --no-e2e"
$CMD3 > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gemv_colmajor -s 12 -c 2 -o gpurun_out/r02/c09_full_c3 $CMD3 > gpurun_out/r02/c09_ncu_full_c3.log 2>&1
BENCH_GRID=128 $CMD > /dev/null 2>&1 && BENCH_GRID=128 ncu --set full --clock-control none --import-source on -k regex:spmv_csr_scalar -s 12 -c 2 -o gpurun_out/r02/c09_full_g128 $CMD > gpurun_out/r02/c09_ncu_full_g128.log 2>&1
ls -la gpurun_out/r02/ | grep c09
