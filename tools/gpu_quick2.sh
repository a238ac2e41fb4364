#!/bin/bash
set -u
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_at_size.py tests/test_gpu_fused_convc1.py -x -q -m gpu -k "${1:-tiled or vol16 or golden or corrblock or full_size or fused or layouts}" 2>&1 | tail -3
show() { python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['value'],1), round(d['ms_per_step'],4), {k:round(v['ms'],4) for k,v in d['kernels'].items()}, 'whole-step frac', round(d['roofline']['whole_step']['frac'],3), {k:(round(v['ms_per_step'],4), round(v.get('fused_call_us',0),1)) for k,v in d.get('fused',{}).items() if isinstance(v,dict)}, {k:round(v['value']) for k,v in d.get('fast_paths',{}).items()})"; }
python bench.py --steps 50 --no-cpu-baseline --no-e2e --no-gpu-reference | show x3
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active
ncu --metrics $M --clock-control none -k regex:lookup_tiled_fwd_kernel -s 40 -c 1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-paths --no-gpu-reference 2>&1 | grep -E "gpu__time|smsp__"
