#!/bin/bash
# round 2, GPU call 11 (two B200s): multi-GPU block power iteration (group + process-per-GPU), group tests after the
# persistent-worker change, warm phase trace of the group one-shot call
mkdir -p gpurun_out/r02
timeout 900 python -m pytest tests/test_local_group.py tests/test_dropin_cpp.py -m gpu -x -q > gpurun_out/r02/c11_pytest_group.log 2>&1; echo "rc=$?" >> gpurun_out/r02/c11_pytest_group.log
tail -12 gpurun_out/r02/c11_pytest_group.log
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/r02/c11_pytest_mgpu.log 2>&1; echo "rc=$?" >> gpurun_out/r02/c11_pytest_mgpu.log
tail -12 gpurun_out/r02/c11_pytest_mgpu.log
EIGSOL_B200_NGPU=2 EIGSOL_B200_TRACE=1 timeout 600 python tools/r02/e2e_trace.py > gpurun_out/r02/c11_e2e_trace_n2.log 2>&1
grep -E "== call" gpurun_out/r02/c11_e2e_trace_n2.log
awk '/== call 1/{f=1} f' gpurun_out/r02/c11_e2e_trace_n2.log | tail -45
for t in 2 4 6 8; do EIGSOL_B200_UPLOAD_THREADS=$t EIGSOL_B200_NGPU=2 timeout 300 python tools/r02/e2e_trace.py 2>&1 | grep "== call 2" | sed "s/^/N=2 threads-per-rank $t /"; done | tee gpurun_out/r02/c11_group_threads.txt
