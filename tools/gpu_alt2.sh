#!/bin/bash
set -u
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_at_size.py tests/test_gpu_raft_forward.py -x -q -m gpu -k "alt or alternate" 2>&1 | tail -6
timeout 300 python tools/bench_configs.py alt4k altsmall 2>&1 | tail -2
RAFT_CORR_ALT_CTAS=1 timeout 300 python tools/bench_configs.py alt4k altsmall 2>&1 | tail -2
