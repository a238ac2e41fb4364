#!/bin/bash
# multi-GPU pass: usage gpu_r2_multi.sh "<probe Ns>" "<bench Ns>" "<ddp Ns>" "<alt-rows Ns>"
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
port=29600
for N in $1; do port=$((port+1))
  timeout 300 $TR --nproc-per-node $N --master-port $port tools/probe_d2h.py > gpurun_out/r02_probe_d2h_n$N.json 2> gpurun_out/probe_n$N.err; echo "probe N=$N rc=$?"; cat gpurun_out/r02_probe_d2h_n$N.json | cut -c1-1500
done
for N in $2; do port=$((port+1))
  timeout 600 $TR --nproc-per-node $N --master-port $port bench.py --gpus $N --steps 30 --warmup 3 > gpurun_out/r02_bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench N=$N rc=$?"
  python -c "import json; d=json.loads(open('gpurun_out/r02_bench_n$N.json').read().strip().splitlines()[-1]); print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d.get('e2e'))"; tail -2 gpurun_out/bench_n$N.err
done
for N in $3; do port=$((port+1))
  timeout 300 $TR --nproc-per-node $N --master-port $port tools/train_step_ddp.py --steps 20 > gpurun_out/r02_ddp_train_n$N.json 2> gpurun_out/ddp_n$N.err; echo "ddp N=$N rc=$?"; cat gpurun_out/r02_ddp_train_n$N.json; tail -2 gpurun_out/ddp_n$N.err
done
for N in $4; do port=$((port+1))
  timeout 300 $TR --nproc-per-node $N --master-port $port tools/alt_rows_sharded.py --check > gpurun_out/r02_alt_rows_n$N.json 2> gpurun_out/altrows_n$N.err; echo "alt rows N=$N rc=$?"; cat gpurun_out/r02_alt_rows_n$N.json; tail -2 gpurun_out/altrows_n$N.err
done
