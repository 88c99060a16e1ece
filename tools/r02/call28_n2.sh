#!/bin/bash
# round 2, GPU call 28 (two B200s): the e2e child alone vs the trace script (same call, same box)
for rep in 1 2; do
python bench.py --e2e-child 2 --workload c5 --steps 20 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('child alone: ', round(d['seconds']*1e3,1), 'ms')"
EIGSOL_B200_NGPU=2 python tools/r02/e2e_trace.py 2>&1 | grep "== call [12]"
done
EIGSOL_B200_TRACE=1 python bench.py --e2e-child 2 --workload c5 --steps 20 2>&1 | grep -E "slab upload|initialised|end \(|iterate" | tail -8
