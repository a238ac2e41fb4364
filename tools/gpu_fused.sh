#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fused_convc1.py -q -m gpu 2>&1 | tail -8
timeout 600 python bench.py --steps 30 --no-cpu-baseline --no-e2e --no-gpu-reference > gpurun_out/bench_fused.json 2> gpurun_out/bench_fused.err; echo "bench rc=$?"
python -c "import json; d=json.load(open('gpurun_out/bench_fused.json')); print(json.dumps(d.get('fused'), indent=1)); print(d['value'], d['kernels'])"
tail -3 gpurun_out/bench_fused.err
