#!/bin/bash
# round 2, GPU call 22 (two B200s): whole-batch graphs in eigsol_b200_power_step -- parity (stepwise API, both process
# models) and the driver-protocol numbers at N=1 and N=2
mkdir -p gpurun_out/r02
timeout 900 python -m pytest tests/test_local_group.py tests/test_multi_gpu.py tests/test_gpu_parity.py -m gpu -x -q -k "not full_size" > gpurun_out/r02/c22_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r02/c22_pytest.log
tail -3 gpurun_out/r02/c22_pytest.log
for rep in 1 2; do
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('c5 N=1 20/5', d['value'], d['ms_per_step'], d['gpu_launches'])"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('c5 N=2 20/5', d['value'], d['ms_per_step'], d['gpu_launches'])"
done
EIGSOL_B200_GRAPH=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('c5 N=2 20/5 no graphs', d['value'], d['ms_per_step'])"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 200 --warmup 20 --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('c5 N=2 200/20', d['value'], d['ms_per_step'])"
for w in c2 c4; do timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline --no-e2e | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$w N=1 20/5', d['value'], d['ms_per_step'])"; done
