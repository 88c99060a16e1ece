timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu19.log 2>&1; echo "rc=$?"; tail -12 gpurun_out/pytest_gpu19.log
python tools/variant_bench.py c2 c4 c5 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l)
    if d['kernel'] in ('csr_window','csr_scalar','sell'): print(d['cfg'], d['kernel'], d['lanes'], 'us=%.1f'%(d['ms_per_trip']*1e3), 'GBps=%.0f'%d['alg_GBps'], 'set=%.2fs'%d['set_kernel_s'])"
