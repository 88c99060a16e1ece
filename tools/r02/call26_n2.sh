#!/bin/bash
# round 2, GPU call 26 (two B200s): does torch / NCCL in the process slow the group upload down?
for mode in "" cuda nccl; do
  TRACE_TORCH=$mode EIGSOL_B200_NGPU=2 timeout 300 python tools/r02/e2e_trace.py 2>&1 | grep "== call [12]" | sed "s/^/torch=[$mode] /"
done
OMP_NUM_THREADS=1 EIGSOL_B200_NGPU=2 timeout 300 python tools/r02/e2e_trace.py 2>&1 | grep "== call [12]" | sed "s/^/OMP_NUM_THREADS=1 /"
