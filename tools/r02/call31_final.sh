#!/bin/bash
# round 2, GPU call 31 (one B200): final state — GPU test suite, smoke(), bench line of the shipped tree
mkdir -p gpurun_out/r02
timeout 500 python -m pytest tests -m gpu -x -q > gpurun_out/r02/c31_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02/c31_pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02/c31_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r02/c31_smoke.log
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/r02/c31_bench_c5_n1.json 2> gpurun_out/r02/c31_bench.err; echo "bench rc=$?"
