#!/bin/bash
# timing experiments with alternative builds of the library: gpu_exp.sh "<-D flags A>" "<-D flags B>" ...
set -u
show() { python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['value'],1), round(d['ms_per_step'],4), {k:round(v['ms'],4) for k,v in d['kernels'].items()})"; }
ARGS="--steps 50 --no-cpu-baseline --no-e2e --no-fast-paths --no-gpu-reference ${BENCH_ARGS:-}"
python bench.py $ARGS | show base
i=0
for flags in "$@"; do i=$((i+1))
  (cd raft_ros_b200 && /usr/local/cuda/bin/nvcc -ccbin /usr/bin/g++ -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC $flags -shared -o /tmp/libexp$i.so csrc/*.cu) || { echo "build failed: $flags"; continue; }
  RAFTCORR_LIB=/tmp/libexp$i.so python bench.py $ARGS | show "[$flags]"
done
python bench.py $ARGS | show base
