#!/bin/bash
# usage: gpu_prof_kernel.sh <kernel regex> <out name> [extra bench args]
set -u
mkdir -p gpurun_out
K=$1; OUT=$2; shift 2
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-paths $*"
$CMD > gpurun_out/plain_$OUT.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -o gpurun_out/$OUT -f $CMD > gpurun_out/ncu_$OUT.log 2>&1; echo "ncu rc=$?"
