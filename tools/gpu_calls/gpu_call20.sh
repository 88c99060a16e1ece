set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu20.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/pytest_gpu20.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke20.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/smoke20.log
python bench.py > gpurun_out/bench20_c5.json 2> gpurun_out/bench20_c5.err; echo "rc=$?"
python bench.py --impl reference --steps 5 > gpurun_out/bench20_ref_c5.json 2> gpurun_out/bench20_ref_c5.err; echo "rc=$?"
for w in c2 c3 c4; do python bench.py --workload $w --no-cpu-baseline > gpurun_out/bench20_$w.json 2> gpurun_out/bench20_$w.err; done
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches20_c5.csv $B > gpurun_out/ncu_launches.log 2>&1
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:spmv -s 4 -c 2 -o gpurun_out/prof20_c5 $B > gpurun_out/ncu_full.log 2>&1
for w in c3 c4; do ncu --set full --clock-control none -k regex:"spmv|gemv" -s 4 -c 1 -o gpurun_out/prof20_$w $B --workload $w > gpurun_out/ncu_full_$w.log 2>&1; done
cat gpurun_out/bench20_c5.json gpurun_out/bench20_ref_c5.json gpurun_out/bench20_c2.json gpurun_out/bench20_c3.json gpurun_out/bench20_c4.json | cut -c1-700
