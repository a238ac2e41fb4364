#!/bin/bash
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/build.log 2>&1; echo "build rc=$?"
timeout 600 python tools/epe_drift.py > gpurun_out/epe_drift.json 2> gpurun_out/epe_drift.err; echo "epe rc=$?"
cat gpurun_out/epe_drift.json; tail -5 gpurun_out/epe_drift.err
