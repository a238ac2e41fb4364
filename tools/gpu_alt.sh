#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_at_size.py tests/test_gpu_raft_forward.py -x -q -m gpu -k "alt or raft_forward or alternate" 2>&1 | tail -8
timeout 300 python tools/bench_configs.py alt4k altsmall 2>&1 | tail -3
