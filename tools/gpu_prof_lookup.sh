#!/bin/bash
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:lookup_fwd -s 36 -c 1 -o gpurun_out/prof_lookup -f $CMD > gpurun_out/ncu_lookup.log 2>&1; echo "ncu lookup rc=$?"
