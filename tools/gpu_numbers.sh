#!/bin/bash
# the numbers the docs quote, from the current code: default bench line, reference arm, all configs, training breakdown
set -u
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r02_bench_n1.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -2 gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_ref.json 2>/dev/null; echo "ref rc=$?"
timeout 300 python bench.py --precision bf16 --no-cpu-baseline --no-e2e --no-gpu-reference --no-fast-paths > gpurun_out/r02_bench_n1_bf16.json 2>/dev/null; echo "bf16 rc=$?"
timeout 900 python tools/bench_configs.py headline sintel32 kitti train altsmall alt4k > gpurun_out/r02_configs.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"; cat gpurun_out/r02_configs.jsonl
timeout 300 python tools/train_breakdown.py > gpurun_out/r02_train_breakdown.json 2>/dev/null; cat gpurun_out/r02_train_breakdown.json
timeout 300 python tools/train_step_ddp.py --steps 20 > gpurun_out/r02_ddp_train_n1.json 2>/dev/null; cat gpurun_out/r02_ddp_train_n1.json
timeout 300 python tools/alt_rows_sharded.py > gpurun_out/r02_alt_rows_n1.json 2>/dev/null; cat gpurun_out/r02_alt_rows_n1.json
timeout 300 python tools/epe_drift.py > gpurun_out/r02_epe_drift.json 2> gpurun_out/epe.err; echo "epe rc=$?"; tail -c 600 gpurun_out/r02_epe_drift.json
