#!/bin/bash
# round 2, GPU call 7 (eight B200s): the driver's scaling protocol (N = 1, 2, 4, 8 at --steps 20 --warmup 5), the same at
# 200 / 20, multi-GPU parity tests at world 8 (process-per-GPU and single-process group), where a trip waits (peer stats)
mkdir -p gpurun_out/r02
nvidia-smi -L > gpurun_out/r02/c07_gpus.txt; nproc >> gpurun_out/r02/c07_gpus.txt
run_bench() {  # N steps warmup tag
  if [ $1 -eq 1 ]; then timeout 600 python bench.py --gpus 1 --steps $2 --warmup $3 > gpurun_out/r02/c07_bench_c5_n$1_$4.json 2> gpurun_out/r02/c07_bench_n$1_$4.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29700 + $1)) bench.py --gpus $1 --steps $2 --warmup $3 > gpurun_out/r02/c07_bench_c5_n$1_$4.json 2> gpurun_out/r02/c07_bench_n$1_$4.err; fi
  python - gpurun_out/r02/c07_bench_c5_n$1_$4.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], 'value', round(d['value'],1), 'us/trip', round(d['ms_per_step']*1e3,2), 'e2e', round(d['e2e']['value'],1) if d.get('e2e') else None, 'parity', d.get('parity') and (d['parity']['lambda_rel_err'], d['parity']['ok']))
except Exception as e:
    print(sys.argv[1], 'FAILED', e)
PY
}
for n in 1 2 4 8; do run_bench $n 20 5 drv; done
for n in 8 4 2; do run_bench $n 200 20 long; done
tail -3 gpurun_out/r02/c07_bench_n8_drv.err
EIGSOL_B200_PEER_STATS=1 LAT_PER_RANK=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29750 tools/latency_bench.py 256 > gpurun_out/r02/c07_peer_stats_n8.txt 2>&1
grep -E "peer_stats|us_per_trip" gpurun_out/r02/c07_peer_stats_n8.txt | head -12
timeout 600 python -m pytest tests/test_local_group.py tests/test_multi_gpu.py tests/test_dropin_cpp.py -m gpu -q -rs -k "8" > gpurun_out/r02/c07_pytest_mgpu8.log 2>&1; echo "rc=$?" >> gpurun_out/r02/c07_pytest_mgpu8.log
tail -6 gpurun_out/r02/c07_pytest_mgpu8.log
for w in c3 c4; do timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29760 bench.py --gpus 8 --steps 200 --warmup 20 --workload $w > gpurun_out/r02/c07_bench_${w}_n8.json 2>/dev/null; done
