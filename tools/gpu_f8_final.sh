#!/bin/bash
# the fp16_fp8 headline: full GPU suite of the build, the bench line, and one ncu capture of the build kernel
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_at_size.py tests/test_gpu_fused_convc1.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -2
python bench.py --steps 50 > gpurun_out/r02_bench_n1.json 2> gpurun_out/bench.err; python -c "
import json; d=json.loads(open('gpurun_out/r02_bench_n1.json').read().strip().splitlines()[-1])
print(d['precision'], 'value', round(d['value'],1), 'ms', round(d['ms_per_step'],4), 'frac', round(d['roofline']['whole_step']['frac'],4), {k:round(v['ms'],4) for k,v in d['kernels'].items()})
print('fast', {k: round(v['value']) for k, v in d.get('fast_paths', {}).items()}, 'e2e', round(d['e2e']['value'],1), 'traffic', d['roofline'].get('traffic'))"
bash tools/gpu_prof_any.sh build_tc_kernel r02_build_tc_fp16_fp8 3 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-paths --no-gpu-reference --precision fp16_fp8 | tail -1 | cut -c1-100
M=gpu__time_duration.sum
ncu --metrics $M --clock-control none -c 60 --csv --log-file gpurun_out/r02_launches_fp16_fp8.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-paths --no-gpu-reference > /dev/null 2>&1; tail -3 gpurun_out/r02_launches_fp16_fp8.csv | cut -c1-200
