timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu9.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/pytest_gpu9.log
for serp in 1 0; do for pin in 0 24 40 50 64 80 100; do
echo "serp=$serp pin=$pin"
EIGSOL_B200_SERPENTINE=$serp EIGSOL_B200_L2_PIN_MB=$pin python tools/latency_bench.py 128 256 2>&1 | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('   g=%d us=%.2f GBps=%.0f'%(d['g'],d['us_per_trip'],d['alg_GBps_per_rank']))"
done; done | tee gpurun_out/l2pin_sweep.txt
