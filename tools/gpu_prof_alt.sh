#!/bin/bash
set -u
mkdir -p gpurun_out
CMD="python tools/bench_configs.py alt4k"
$CMD > gpurun_out/plain_alt.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:alt_tc_kernel -s 2 -c 1 -o gpurun_out/prof_alt -f $CMD > gpurun_out/ncu_alt.log 2>&1; echo "ncu rc=$?"
cat gpurun_out/plain_alt.log | cut -c1-300
