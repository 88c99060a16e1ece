timeout 400 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/pytest_gpu46.log 2>&1; echo "gpu suite rc=$?"; tail -4 gpurun_out/pytest_gpu46.log | cut -c1-200
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 300 python bench.py > gpurun_out/bench46_c5.json 2> gpurun_out/bench46_c5.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench46_c5.json')); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e'], d['cpu_baseline'], d['gpu_launches'], d['clocks'])"
