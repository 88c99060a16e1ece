python tools/latency_bench.py 203 128 2>&1 | grep '^{' | cut -c1-120
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/latency_bench.py 128 256"
echo default; timeout 200 $T 2>/dev/null | grep '^{' | cut -c1-120
echo nodefer; EIGSOL_B200_DEFER=0 timeout 200 $T 2>/dev/null | grep '^{' | cut -c1-120
echo nosend; EIGSOL_B200_DEBUG_NOSEND=1 timeout 200 $T 2>/dev/null | grep '^{' | cut -c1-120
echo nosend-nodefer; EIGSOL_B200_DEFER=0 EIGSOL_B200_DEBUG_NOSEND=1 timeout 200 $T 2>/dev/null | grep '^{' | cut -c1-120
echo noserp; EIGSOL_B200_SERPENTINE=0 timeout 200 $T 2>/dev/null | grep '^{' | cut -c1-120
