#!/bin/bash
set -u
timeout 600 python -m pytest tests/test_gpu_fused_convc1.py tests/test_gpu_encoder_head.py -x -q -m gpu 2>&1 | tail -12
timeout 300 python bench.py --steps 30 --no-cpu-baseline --no-e2e --no-gpu-reference 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(json.dumps(d.get('fused'), indent=1)); print(d['value'], d['ms_per_step'])"
