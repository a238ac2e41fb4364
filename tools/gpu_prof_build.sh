#!/bin/bash
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --precision ${1:-bf16x3}"
$CMD > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:build_tc_kernel -s 3 -c 1 -o gpurun_out/prof_build -f $CMD > gpurun_out/ncu_build.log 2>&1; echo "ncu build rc=$?"
