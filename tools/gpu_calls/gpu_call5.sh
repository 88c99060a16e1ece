set -x
nvidia-smi -L
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?"; tail -8 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "rc=$?"
python bench.py --impl reference --steps 5 > gpurun_out/bench_ref_c5.json 2> gpurun_out/bench_ref_c5.err; echo "rc=$?"
for w in c2 c3 c4; do python bench.py --workload $w --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; done
python tools/variant_bench.py > gpurun_out/variants.jsonl 2> gpurun_out/variants.err
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_c5.csv $B > gpurun_out/ncu_launches.log 2>&1
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:spmv -s 4 -c 2 -o gpurun_out/prof_c5 $B > gpurun_out/ncu_full.log 2>&1
cat gpurun_out/bench_c5.json; cat gpurun_out/bench_ref_c5.json; tail -3 gpurun_out/bench_c5.err
cat gpurun_out/bench_c2.json gpurun_out/bench_c3.json gpurun_out/bench_c4.json
