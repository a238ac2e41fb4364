#!/bin/bash
set -u
mkdir -p gpurun_out
CMD="python tools/bench_configs.py train"
$CMD > gpurun_out/plain_train.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_train.csv $CMD > gpurun_out/ncu_train.log 2>&1; echo "ncu rc=$?"
