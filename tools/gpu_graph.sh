#!/bin/bash
set -u
timeout 600 python -m pytest tests/test_gpu_cuda_graph.py -q -m gpu 2>&1 | tail -8
timeout 300 python tools/bench_latency.py 2>&1 | tail -3
