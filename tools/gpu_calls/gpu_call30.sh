python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu30.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/pytest_gpu30.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py > gpurun_out/bench30_c5.json 2> gpurun_out/bench30_c5.err; echo "rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench30_c5.json')); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['traffic'], d['e2e']['value'], d['cpu_baseline']['value'], d['clocks'])"
