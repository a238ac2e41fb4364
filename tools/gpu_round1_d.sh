#!/bin/bash
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/build.log 2>&1; echo "build rc=$?"
timeout 600 python -m pytest tests -m gpu -q -k "reference_binary" > gpurun_out/pytest_refbin.log 2>&1; echo "refbin rc=$?"; tail -3 gpurun_out/pytest_refbin.log
timeout 900 python tools/bench_configs.py headline kitti train altsmall alt4k > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"
cat gpurun_out/configs.jsonl; tail -3 gpurun_out/configs.err
