#!/bin/bash
# round 2, GPU call 2: A/B of the round-1 library (build/r01tree, commit 8f8aa0f) against the current one on the SAME box
mkdir -p gpurun_out/r02
for w in c5 c4 c3 c2; do
  for rep in 1 2; do
    (cd build/r01tree && timeout 300 python bench.py --workload $w --steps 200 --warmup 20 --no-cpu-baseline --no-e2e) 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('r01 $w', round(d['ms_per_step']*1e3,2),'us', round(d['roofline']['frac'],4))"
    timeout 300 python bench.py --workload $w --steps 200 --warmup 20 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('r02 $w', round(d['ms_per_step']*1e3,2),'us', round(d['roofline']['frac'],4), 'readmix', d['roofline'].get('peak_read_mix'))"
  done
done 2>&1 | tee gpurun_out/r02/c02_ab.txt
python - <<'PY' 2>&1 | tee gpurun_out/r02/c02_probe.txt
import pcsc_eigenvalue_solver_project_b200 as es
for pb in (128 << 20, 150_000_000 // 4096 * 4096, 134217728 + 65536 * 3):
    for planes in (1, 12):
        print('hbm_probe plane_bytes', pb, 'planes', planes, es.hbm_probe(pb, planes, 30))
import torch
a = torch.empty(1 << 30, dtype=torch.bfloat16, device='cuda'); b = torch.empty_like(a)
best = 0
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); b.copy_(a); e1.record(); torch.cuda.synchronize()
    best = max(best, 2 * a.numel() * 2 / (e0.elapsed_time(e1) * 1e-3) / 1e9)
print('torch copy GB/s (the MEASURED_PEAKS recipe) on this box', best)
PY
nvidia-smi --query-gpu=clocks.sm,clocks.mem,clocks.max.mem,power.draw,temperature.gpu --format=csv >> gpurun_out/r02/c02_probe.txt
