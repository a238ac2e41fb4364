#!/bin/bash
# conv2d_act / conv2d_fused / UpdateBlock / FusedRAFT: parity, per-layer timing against cuDNN, and the whole RAFT.forward
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_update.py -x -q -m gpu -s 2>&1 | grep -v "^$" | tail -12
timeout 600 python tools/bench_conv.py > gpurun_out/conv.jsonl 2> gpurun_out/conv.err; tail -14 gpurun_out/conv.jsonl
timeout 900 python tools/raft_e2e.py > gpurun_out/raft_e2e.json 2> gpurun_out/raft_e2e.err; tail -1 gpurun_out/raft_e2e.json; tail -3 gpurun_out/raft_e2e.err
