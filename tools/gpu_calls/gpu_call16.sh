timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu16.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/pytest_gpu16.log
LAT_CONFIGS="pdl:;nopdl:EIGSOL_B200_PDL=0" python tools/latency_bench.py 16 64 128 256 2>&1 | grep '^{' | cut -c1-140
python tools/variant_bench.py > gpurun_out/variants16.jsonl 2> gpurun_out/variants16.err
python - <<'PY'
import json
for l in open('gpurun_out/variants16.jsonl'):
    d=json.loads(l); print(d['cfg'], d['kernel'], d['lanes'], 'us=%.1f'%(d['ms_per_trip']*1e3), 'GBps=%.0f'%d['alg_GBps'])
PY
