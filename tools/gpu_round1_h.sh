#!/bin/bash
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/build.log 2>&1; echo "build rc=$?"
timeout 300 python -m pytest tests -m gpu -q -k "alternate_block" > gpurun_out/pytest_alt.log 2>&1; echo "pytest alt rc=$?"; tail -25 gpurun_out/pytest_alt.log
timeout 600 python tools/bench_configs.py altsmall alt4k > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"
cat gpurun_out/configs.jsonl; tail -3 gpurun_out/configs.err
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
