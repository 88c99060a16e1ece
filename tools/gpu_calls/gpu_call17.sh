timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu17.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/pytest_gpu17.log
python bench.py --no-cpu-baseline > gpurun_out/bench17_c5.json 2> gpurun_out/bench17_c5.err; python -c "
import json; d=json.load(open('gpurun_out/bench17_c5.json')); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e'])"
tail -3 gpurun_out/bench17_c5.err
for t in 1 2 8; do EIGSOL_B200_UPLOAD_THREADS=$t python bench.py --no-cpu-baseline --steps 50 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('threads $t', d['e2e']['seconds'])"; done
EIGSOL_B200_DEVICE_INGEST=0 python bench.py --no-cpu-baseline --steps 50 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('host ingest', d['e2e']['seconds'])"
