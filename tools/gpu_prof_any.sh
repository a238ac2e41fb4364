#!/bin/bash
# usage: gpu_prof_any.sh <kernel regex> <out name> <skip> <python script + args...>: plain run, then one --set full capture
set -u
mkdir -p gpurun_out
K=$1; OUT=$2; SKIP=$3; shift 3
"$@" > gpurun_out/plain_$OUT.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_$OUT.log; exit 1; }
tail -2 gpurun_out/plain_$OUT.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$K -s $SKIP -c 1 -o gpurun_out/$OUT -f "$@" > gpurun_out/ncu_$OUT.log 2>&1; echo "ncu rc=$?"
