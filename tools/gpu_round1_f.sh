#!/bin/bash
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/build.log 2>&1; echo "build rc=$?"
timeout 300 python -m pytest tests -m gpu -x -q -k "bf16 or no_alloc" > gpurun_out/pytest_tc2.log 2>&1; echo "pytest 2cta rc=$?"; tail -15 gpurun_out/pytest_tc2.log
for n in 2 1; do
  export RAFT_CORR_TC_CTAS=$n
  for p in bf16x3 bf16; do
    timeout 300 python bench.py --precision $p --no-cpu-baseline --no-e2e > gpurun_out/bench_${p}_cta$n.json 2> gpurun_out/bench_${p}_cta$n.err; echo "bench $p cta$n rc=$?"
    python - <<PY
import json
try:
    d=json.load(open('gpurun_out/bench_${p}_cta$n.json'))
    print('$p cta$n', round(d['value']), 'build ms', round(d['kernels']['build']['ms'],4), 'lookup ms', round(d['kernels']['lookup']['ms'],4), d['clocks'])
except Exception as e: print('no result', e)
PY
  done
done
unset RAFT_CORR_TC_CTAS
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest all rc=$?"; tail -4 gpurun_out/pytest_gpu.log
