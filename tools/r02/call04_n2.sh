#!/bin/bash
# round 2, GPU call 4 (two B200s): single-process group tests, process-per-GPU tests, drop-in binaries on a group,
# bench at N=2 under the driver's protocol and at 200/20
mkdir -p gpurun_out/r02
nvidia-smi -L > gpurun_out/r02/c04_gpus.txt; nvidia-smi topo -m >> gpurun_out/r02/c04_gpus.txt 2>&1
timeout 900 python -m pytest tests/test_local_group.py tests/test_dropin_cpp.py -m gpu -x -q -rs > gpurun_out/r02/c04_pytest_group.log 2>&1; echo "rc=$?" >> gpurun_out/r02/c04_pytest_group.log
tail -15 gpurun_out/r02/c04_pytest_group.log
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q -rs > gpurun_out/r02/c04_pytest_mgpu.log 2>&1; echo "rc=$?" >> gpurun_out/r02/c04_pytest_mgpu.log
tail -8 gpurun_out/r02/c04_pytest_mgpu.log
for sw in "20 5" "200 20"; do set -- $sw
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps $1 --warmup $2 > gpurun_out/r02/c04_bench_c5_n2_$1.json 2> gpurun_out/r02/c04_bench_n2_$1.err; echo "bench rc=$?"
  cat gpurun_out/r02/c04_bench_c5_n2_$1.json; tail -5 gpurun_out/r02/c04_bench_n2_$1.err
done
EIGSOL_B200_NGPU=2 EIGSOL_B200_TRACE=1 timeout 600 python tools/r02/e2e_trace.py > gpurun_out/r02/c04_e2e_trace_n2.log 2>&1
grep -E "== call|rank|group" gpurun_out/r02/c04_e2e_trace_n2.log | tail -60
# ---- single-GPU extras riding on this lease ----
one() {  # workload env...
  w=$1; shift
  env "$@" timeout 300 python bench.py --workload $w --steps 200 --warmup 20 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w $*', round(d['ms_per_step']*1e3,2),'us', round(d['roofline']['frac'],4))"
}
{
for wv in 4 5; do one c3 EIGSOL_B200_GEMV_VAR=2 EIGSOL_B200_GEMV_WAVES=$wv; one c3 EIGSOL_B200_GEMV_VAR=2 EIGSOL_B200_GEMV_MAP=1 EIGSOL_B200_GEMV_WAVES=$wv; done
one c3 EIGSOL_B200_GEMV_VAR=2 EIGSOL_B200_GEMV_WAVES=8
one c3 EIGSOL_B200_GEMV_VAR=0 EIGSOL_B200_GEMV_WAVES=5
BENCH_GRID=128 one c5 EIGSOL_B200_SCALAR_BLOCK=256; BENCH_GRID=128 one c5 EIGSOL_B200_SCALAR_BLOCK=128
one c5 EIGSOL_B200_SCALAR_BLOCK=128; one c2 EIGSOL_B200_SCALAR_BLOCK=128
} 2>&1 | tee gpurun_out/r02/c04_extras.txt
for blk in 256 128; do
EIGSOL_B200_SCALAR_BLOCK=$blk timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus 2 --steps 200 --warmup 20 --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N=2 SCALAR_BLOCK=$blk', round(d['ms_per_step']*1e3,2),'us')" | tee -a gpurun_out/r02/c04_extras.txt
done
EIGSOL_B200_PEER_STATS=1 LAT_PER_RANK=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29614 tools/latency_bench.py 128 256 > gpurun_out/r02/c04_peer_stats_n2.txt 2>&1
grep -E "peer_stats|us_per_trip" gpurun_out/r02/c04_peer_stats_n2.txt | head
TRIPS=100 timeout 600 python tools/variant_bench.py s5 s6 s7 s9 s10 > gpurun_out/r02/c04_variants_s.jsonl 2> /dev/null
