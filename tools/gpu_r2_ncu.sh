#!/bin/bash
# round-2 ncu captures (one launch each, --set full): headline kernels, on-the-fly kernel at 4K in both
# precisions, training kernels at config 4.  Each capture only after the same command ran clean without ncu.
set -u
mkdir -p gpurun_out
cap() {  # <out> <kernel regex> <skip> <cmd...>
  local out=$1 k=$2 skip=$3; shift 3
  "$@" > gpurun_out/plain_$out.log 2>&1 || { echo "plain run of $out failed"; tail -3 gpurun_out/plain_$out.log; return; }
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 -o gpurun_out/$out -f "$@" > gpurun_out/ncu_$out.log 2>&1
  echo "ncu $out rc=$?"
}
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-paths --no-gpu-reference"
cap r02_lookup_tiled_fwd lookup_tiled_fwd_kernel 40 $B
cap r02_build_tc_tiled_bf16x3 build_tc_kernel 3 $B
cap r02_alt_tc_bf16x3 alt_tc_kernel 2 python tools/bench_configs.py alt4k_x3
cap r02_alt_tc_bf16 alt_tc_kernel 2 python tools/bench_configs.py alt4k_bf16
cap r02_lookup_bwd lookup_bwd_kernel 14 python tools/bench_configs.py train
cap r02_fold_split fold_split_kernel 1 python tools/bench_configs.py train
cap r02_gemm_tc gemm_tc_kernel 1 python tools/bench_configs.py train
# launch list of the bench step (shares of the step)
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches_bench_steps2.csv $B > gpurun_out/ncu_launches.log 2>&1; echo "launch list rc=$?"
ls -la gpurun_out/*.ncu-rep
