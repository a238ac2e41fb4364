#!/bin/bash
# what the driver's scaling step runs: both arms at N = 1, 2, 4, 8 back to back (needs gpurun --gpus 8)
set -u
mkdir -p gpurun_out
port=29700
for N in 1 2 4 8; do
  port=$((port+2))
  if [ $N -eq 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port"; fi
  timeout 300 $L bench.py --impl reference --gpus $N --steps 5 --warmup 1 2> gpurun_out/scale_ref_n$N.err | tail -1 > gpurun_out/r02_scale_ref_n$N.json; echo "ref N=$N rc=$?"
  if [ $N -ne 1 ]; then L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((port+1))"; fi
  timeout 600 $L bench.py --gpus $N --steps 30 --warmup 3 2> gpurun_out/scale_n$N.err | tail -1 > gpurun_out/r02_scale_n$N.json; echo "ours N=$N rc=$?"
  python -c "
import json
d=json.load(open('gpurun_out/r02_scale_n$N.json')); r=json.load(open('gpurun_out/r02_scale_ref_n$N.json'))
print('N=$N value', round(d['value'],1), 'ms', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value'],1), 'frac_of_probe', d['e2e'].get('frac_of_probe'), 'ref', round(r['value'],2), 'clocks', d['clocks']['sm_mhz'], d['clocks']['reasons'])"
done
