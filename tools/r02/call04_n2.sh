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
