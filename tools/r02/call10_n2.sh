#!/bin/bash
# round 2, GPU call 10 (two B200s): group tests after the shared-pool change, warm phase trace of the group ingest,
# bench N=2 at the driver's protocol, peer stats
mkdir -p gpurun_out/r02
timeout 900 python -m pytest tests/test_local_group.py tests/test_dropin_cpp.py -m gpu -x -q > gpurun_out/r02/c10_pytest_group.log 2>&1; echo "rc=$?" >> gpurun_out/r02/c10_pytest_group.log
tail -4 gpurun_out/r02/c10_pytest_group.log
EIGSOL_B200_NGPU=2 EIGSOL_B200_TRACE=1 timeout 600 python tools/r02/e2e_trace.py > gpurun_out/r02/c10_e2e_trace_n2.log 2>&1
grep -E "== call" gpurun_out/r02/c10_e2e_trace_n2.log
awk '/== call 1/{f=1} f' gpurun_out/r02/c10_e2e_trace_n2.log | tail -50
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02/c10_bench_c5_n2_20.json 2> gpurun_out/r02/c10_bench_n2.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r02/c10_bench_c5_n2_20.json').read().strip().splitlines()[-1]); print('N=2', d['value'], d['ms_per_step'], 'e2e', d['e2e'], 'parity', d['parity'])"
EIGSOL_B200_PEER_STATS=1 LAT_PER_RANK=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29614 tools/latency_bench.py 128 256 > gpurun_out/r02/c10_peer_stats_n2.txt 2>&1
grep -E "peer_stats|us_per_trip" gpurun_out/r02/c10_peer_stats_n2.txt | head
