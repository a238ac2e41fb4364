#!/bin/bash
# round 2, first pass: the new at-size / RAFT.forward parity tests, then the whole GPU suite, smoke, bench
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_at_size.py tests/test_gpu_raft_forward.py -q -m gpu -s > gpurun_out/pytest_new.log 2>&1; echo "new tests rc=$?"
grep -E "raft.forward|passed|failed|Error|assert" gpurun_out/pytest_new.log | head -40
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
head -c 4000 gpurun_out/bench.json; echo
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
head -c 1500 gpurun_out/bench_ref.json; echo
