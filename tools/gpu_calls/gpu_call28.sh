timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu28.log 2>&1; echo "rc=$?"; tail -8 gpurun_out/pytest_gpu28.log | cut -c1-300
EIGSOL_B200_TRACE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 200 --no-cpu-baseline 2>gpurun_out/bench28.err > gpurun_out/bench28_c5_n2.json; grep "eigsol_b200\]" gpurun_out/bench28.err | tail -16; python -c "
import json; d=json.load(open('gpurun_out/bench28_c5_n2.json')); print('c5 n2', d['value'], d['ms_per_step'], d['e2e'])"
python bench.py --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('c5 n1', d['value'], d['ms_per_step'], d['e2e']['value'], d['build_s'])"
