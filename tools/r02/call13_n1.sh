#!/bin/bash
# round 2, GPU call 13 (one B200): ncu evidence (launch list + full captures), staging-copy variants, full GPU suite
mkdir -p gpurun_out/r02
for nt in 0 1; do for rep in 1 2; do EIGSOL_B200_STAGE_NT=$nt timeout 300 python tools/r02/e2e_trace.py 2>&1 | grep "== call [12]" | sed "s/^/STAGE_NT=$nt /"; done; done | tee gpurun_out/r02/c13_stage_nt.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02/c13_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02/c13_pytest.log
tail -5 gpurun_out/r02/c13_pytest.log
bash tools/r02/call09_ncu.sh
