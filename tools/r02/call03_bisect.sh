#!/bin/bash
# round 2, GPU call 3: why is every kernel ~14 % slower than round 1 on the same box?  env toggles on both builds
mkdir -p gpurun_out/r02
run() {  # $1 = tree label, $2 = dir, rest = env assignments
  label=$1; dir=$2; shift 2
  for w in c3 c5; do
    (cd $dir && env "$@" timeout 300 python bench.py --workload $w --steps 200 --warmup 20 --no-cpu-baseline --no-e2e) 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$label $w $*', round(d['ms_per_step']*1e3,2),'us')"
  done
}
{
run r01 build/r01tree X=1
run r02 . X=1
run r01 build/r01tree EIGSOL_B200_GRAPH=0
run r02 . EIGSOL_B200_GRAPH=0
run r01 build/r01tree EIGSOL_B200_PDL=0
run r02 . EIGSOL_B200_PDL=0
run r01 build/r01tree EIGSOL_B200_GRAPH=0 EIGSOL_B200_PDL=0
run r02 . EIGSOL_B200_GRAPH=0 EIGSOL_B200_PDL=0
run r02 . EIGSOL_B200_GRAPH=3
run r02 . EIGSOL_B200_POOL_KEEP_MB=0
} 2>&1 | tee gpurun_out/r02/c03_bisect.txt
