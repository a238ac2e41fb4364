#!/bin/bash
# tiled-layout check: tests, then bench in both layouts
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tiled or relayout or layouts" > gpurun_out/tiled_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/tiled_tests.log
tail -5 gpurun_out/tiled_tests.log
for lay in tiled rows; do
  timeout 600 python bench.py --steps 50 --no-cpu-baseline --layout $lay > gpurun_out/bench_$lay.json 2> gpurun_out/bench_$lay.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/bench_$lay.json").read().strip().splitlines()[-1])
print("$lay", round(d["value"],1), "ms", round(d["ms_per_step"],4), {k:round(v["ms"],4) for k,v in d["kernels"].items()}, "e2e", round(d["e2e"]["value"],1))
PY
done
for lay in tiled rows; do
  timeout 600 python bench.py --steps 30 --no-cpu-baseline --no-e2e --layout $lay --height 376 --width 1248 > gpurun_out/bench_kitti_$lay.json 2> gpurun_out/bench_kitti_$lay.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/bench_kitti_$lay.json").read().strip().splitlines()[-1])
print("kitti $lay", round(d["value"],1), "ms", round(d["ms_per_step"],4), {k:round(v["ms"],4) for k,v in d["kernels"].items()})
PY
done
