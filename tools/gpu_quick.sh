#!/bin/bash
# quick check after a build-kernel change: build-related parity tests + bench in both precisions
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "${1:-build or golden or tiled or corrblock}" 2>&1 | tail -2
show() { python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['value'],1), round(d['ms_per_step'],4), {k:round(v['ms'],4) for k,v in d['kernels'].items()}, 'whole-step frac', round(d['roofline']['whole_step']['frac'],3))"; }
python bench.py --steps 50 --no-cpu-baseline --no-e2e | show x3
python bench.py --steps 50 --no-cpu-baseline --no-e2e --precision bf16 | show bf16
