#!/bin/bash
# programmatic dependent launch on/off: parity of the touched kernels, then the bench step both ways
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_raft_forward.py tests/test_zz_no_alloc.py -x -q -m gpu -s 2>&1 | grep -E "raft.forward|passed|failed|rror" | head -20
show() { python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['value'],1), round(d['ms_per_step'],4), {k:round(v['ms'],4) for k,v in d['kernels'].items()}, 'whole-step frac', round(d['roofline']['whole_step']['frac'],3))"; }
for i in 1 2; do
RAFT_CORR_PDL=0 python bench.py --steps 100 --no-cpu-baseline --no-e2e --no-fast-paths | show pdl0
RAFT_CORR_PDL=1 python bench.py --steps 100 --no-cpu-baseline --no-e2e --no-fast-paths | show pdl1
done
RAFT_CORR_PDL=0 python bench.py --steps 100 --no-cpu-baseline --no-e2e --no-fast-paths --precision bf16 | show bf16_pdl0
RAFT_CORR_PDL=1 python bench.py --steps 100 --no-cpu-baseline --no-e2e --no-fast-paths --precision bf16 | show bf16_pdl1
