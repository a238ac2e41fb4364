#!/bin/bash
set -u
timeout 300 python -m pytest tests/test_gpu_upsample.py -q -m gpu 2>&1 | tail -5
python - <<'PY'
import sys, torch, json
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import raft_ros_b200 as rr
from test_gpu_upsample import reference_upsample_flow
g = torch.Generator().manual_seed(1)
N, H, W = 8, 55, 128
flow = torch.randn(N, 2, H, W, generator=g).cuda(); mask = torch.randn(N, 576, H, W, generator=g).cuda()
def t(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); a, e = torch.cuda.Event(True), torch.cuda.Event(True); a.record()
    for _ in range(reps): fn()
    e.record(); torch.cuda.synchronize(); return a.elapsed_time(e) / reps * 1e3
print(json.dumps({"ours_us": t(lambda: rr.upsample_flow(flow, mask)), "reference_torch_ops_us": t(lambda: reference_upsample_flow(flow, mask)),
                  "bytes": N * H * W * (576 + 128 + 2) * 4}))
PY
