set -u
mkdir -p gpurun_out
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711"
timeout 300 $L bench.py --impl reference --gpus 2 --steps 5 --warmup 1 2> gpurun_out/s2r.err | tail -1 | cut -c1-200
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29713"
timeout 600 $L bench.py --gpus 2 --steps 30 --warmup 3 2> gpurun_out/s2.err | tail -1 > gpurun_out/r02_scale_n2.json
python -c "
import json
d=json.load(open('gpurun_out/r02_scale_n2.json'))
print('N=2 value', round(d['value'],1), 'ms', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value'],1), d['e2e'].get('frac_of_probe'), 'keys', sorted(d.keys()))"
timeout 300 python -m pytest tests/test_shard.py -q 2>&1 | tail -1
