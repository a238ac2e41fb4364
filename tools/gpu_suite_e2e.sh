#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
for b in 1 0 1; do
RAFT_BENCH_NUMA_BIND=$b python bench.py --steps 20 --no-cpu-baseline --no-fast-paths --no-gpu-reference 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bind$b', round(d['value'],1), d['e2e']['value'], d['e2e']['d2h_gbs'])"
done
python tools/probe_d2h.py | cut -c1-700
