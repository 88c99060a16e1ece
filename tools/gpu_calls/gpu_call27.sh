for w in c5 c4 c3 c2; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 4 --workload $w $( [ $w != c5 ] && echo --no-e2e ) 2>gpurun_out/bench27_${w}_n4.err > gpurun_out/bench27_${w}_n4.json; python -c "
import json; d=json.load(open('gpurun_out/bench27_${w}_n4.json')); print('$w', d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e'] and d['e2e']['value'])"
done
