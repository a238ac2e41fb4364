#!/bin/bash
# what the driver runs at round end, on one GPU: GPU tests, smoke, reference arm, our arm (timed by wall clock too)
set -u
mkdir -p gpurun_out
t0=$(date +%s); timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$? ($(( $(date +%s) - t0 )) s)"; tail -3 gpurun_out/pytest_gpu.log
t0=$(date +%s); timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$? ($(( $(date +%s) - t0 )) s)"; tail -4 gpurun_out/smoke.log
t0=$(date +%s); timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 3 > gpurun_out/r02_bench_ref.json 2> gpurun_out/bench_ref.err; echo "reference arm rc=$? ($(( $(date +%s) - t0 )) s)"; cut -c1-300 gpurun_out/r02_bench_ref.json
t0=$(date +%s); timeout 900 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/r02_bench_n1.json 2> gpurun_out/bench.err; echo "our arm rc=$? ($(( $(date +%s) - t0 )) s)"
python -c "
import json; d=json.loads(open('gpurun_out/r02_bench_n1.json').read().strip().splitlines()[-1])
print('value', round(d['value'],1), 'ms', round(d['ms_per_step'],4), 'frac', round(d['roofline']['whole_step']['frac'],4), 'dom', d['roofline']['kernel'], round(d['roofline']['frac'],4))
print('e2e', d['e2e']['value'], 'cpu', d['cpu_baseline'], 'launches', d['gpu_launches'])
print('fused', {k: v['ms_per_step'] for k, v in d['fused'].items() if isinstance(v, dict)}, 'gpu_ref', d['gpu_reference']['corrblock']['value'], d['gpu_reference'].get('alt_corr_4k'))
"
