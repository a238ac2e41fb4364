#!/bin/bash
# First GPU pass: parity (non tensor-core first, then tcgen05 under a timeout), smoke, bench, launch list.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/build.log 2>&1; echo "build rc=$?"
timeout 900 python -m pytest tests -m gpu -q -k "not bf16" > gpurun_out/pytest_simt.log 2>&1; echo "pytest(non-tc) rc=$?"
tail -5 gpurun_out/pytest_simt.log
timeout 300 python bench.py --precision fp32 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_fp32.json 2> gpurun_out/bench_fp32.err; echo "bench fp32 rc=$?"
cat gpurun_out/bench_fp32.json | head -c 1500; echo
timeout 600 python -m pytest tests -m gpu -q -k "bf16" > gpurun_out/pytest_tc.log 2>&1; echo "pytest(tc) rc=$?"
tail -30 gpurun_out/pytest_tc.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
tail -5 gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_bf16x3.json 2> gpurun_out/bench_bf16x3.err; echo "bench bf16x3 rc=$?"
cat gpurun_out/bench_bf16x3.json | head -c 3000; echo
timeout 300 python bench.py --precision bf16 --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_bf16.json 2> gpurun_out/bench_bf16.err; echo "bench bf16 rc=$?"
cat gpurun_out/bench_bf16.json | head -c 1500; echo
