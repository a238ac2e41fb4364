#!/bin/bash
set -u
mkdir -p gpurun_out
N=${1:-2}
python tools/train_step_ddp.py --steps 20 > gpurun_out/ddp_n1.json 2> gpurun_out/ddp_n1.err; echo "ddp N=1 rc=$?"; cat gpurun_out/ddp_n1.json; tail -2 gpurun_out/ddp_n1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 tools/train_step_ddp.py --steps 20 > gpurun_out/ddp_n$N.json 2> gpurun_out/ddp_n$N.err; echo "ddp N=$N rc=$?"; cat gpurun_out/ddp_n$N.json; tail -2 gpurun_out/ddp_n$N.err
