#!/bin/bash
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/build.log 2>&1; echo "build rc=$?"
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print(d['value'], d['ms_per_step'], json.dumps(d['kernels']), d['clocks'], d.get('e2e'))
PY
python tools/probe_bw.py > gpurun_out/probe_bw.txt 2>&1; cat gpurun_out/probe_bw.txt
timeout 300 python bench.py --precision bf16 --no-cpu-baseline --no-e2e > gpurun_out/bench_bf16.json 2> gpurun_out/bench_bf16.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_bf16.json'))
print('bf16', d['value'], d['ms_per_step'], json.dumps(d['kernels']))
PY
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
$CMD > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:build_tc_kernel -s 3 -c 1 -o gpurun_out/prof_build -f $CMD > gpurun_out/ncu_build.log 2>&1; echo "ncu build rc=$?"
