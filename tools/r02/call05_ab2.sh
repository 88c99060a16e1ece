#!/bin/bash
# round 2, GPU call 5: A/B against round 1 after restoring the non-coherent gathers for the one-GPU kernels; GEMV tail /
# grid variants; HBM probe variants; e2e trace and upload thread / chunk sweep
mkdir -p gpurun_out/r02
one() {  # label dir workload env...
  label=$1; dir=$2; w=$3; shift 3
  (cd $dir && env "$@" timeout 300 python bench.py --workload $w --steps 200 --warmup 20 --no-cpu-baseline --no-e2e) 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$label $w $*', round(d['ms_per_step']*1e3,2),'us', round(d['roofline']['frac'],4), d['roofline'].get('peak_read_mix'))"
}
{
for w in c5 c4 c3 c2; do one r01 build/r01tree $w X=1; one r02 . $w X=1; done
one r02 . c3 EIGSOL_B200_GEMV_TAIL=1
one r02 . c3 EIGSOL_B200_GEMV_WAVES=5
one r02 . c3 EIGSOL_B200_GEMV_WAVES=5 EIGSOL_B200_GEMV_TAIL=1
one r02 . c3 EIGSOL_B200_GEMV_WAVES=3 EIGSOL_B200_GEMV_TAIL=1
} 2>&1 | tee gpurun_out/r02/c05_ab.txt
python - <<'PY' 2>&1 | tee gpurun_out/r02/c05_probe.txt
import os
import pcsc_eigenvalue_solver_project_b200 as es
for ctas in (2, 4, 8):
    os.environ["EIGSOL_B200_PROBE_CTAS_PER_SM"] = str(ctas)
    for planes in (1, 4, 12):
        print('hbm_probe ctas/SM', ctas, 'planes', planes, es.hbm_probe(150_000_000 // 4096 * 4096, planes, 30))
PY
EIGSOL_B200_TRACE=1 timeout 600 python tools/r02/e2e_trace.py > gpurun_out/r02/c05_e2e_trace.log 2>&1
tail -24 gpurun_out/r02/c05_e2e_trace.log
for cfg in "8 4" "12 4" "16 4" "8 2" "8 8" "16 2" "12 8"; do set -- $cfg
  EIGSOL_B200_UPLOAD_THREADS=$1 EIGSOL_B200_UPLOAD_CHUNK_MB=$2 timeout 300 python tools/r02/e2e_trace.py 2>&1 | grep "== call 2" | sed "s/^/threads $1 chunkMB $2 /"
done | tee gpurun_out/r02/c05_upload_sweep.txt
