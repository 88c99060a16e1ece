set -x
python bench.py > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "rc=$?"
python bench.py --impl reference --steps 5 > gpurun_out/bench_ref_c5.json 2> gpurun_out/bench_ref_c5.err; echo "rc=$?"
for w in c2 c3 c4; do python bench.py --workload $w --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; done
python tools/variant_bench.py s16 s32 s64 i8 i32 > gpurun_out/variants_tune.jsonl 2> gpurun_out/variants_tune.err
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_c5.csv $B > gpurun_out/ncu_launches.log 2>&1
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:spmv -s 4 -c 2 -o gpurun_out/prof_c5 $B > gpurun_out/ncu_full.log 2>&1
cat gpurun_out/bench_c5.json; cat gpurun_out/bench_ref_c5.json; tail -3 gpurun_out/bench_c5.err
