#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tiled or relayout or layouts" 2>&1 | tail -2
show() { python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['value'],1), round(d['ms_per_step'],4), {k:round(v['ms'],4) for k,v in d['kernels'].items()})"; }
python bench.py --steps 50 --no-cpu-baseline --no-e2e | show G32
python bench.py --steps 30 --no-cpu-baseline --no-e2e --height 376 --width 1248 --iters 24 | show G32-kitti
for G in 8 16; do
  RAFTCORR_LIB=$PWD/raft_ros_b200/libraftcorr_expG$G.so python bench.py --steps 50 --no-cpu-baseline --no-e2e | show G$G
  RAFTCORR_LIB=$PWD/raft_ros_b200/libraftcorr_expG$G.so python bench.py --steps 30 --no-cpu-baseline --no-e2e --height 376 --width 1248 --iters 24 | show G$G-kitti
done
