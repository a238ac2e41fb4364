#!/bin/bash
set -u
timeout 600 python -m pytest tests/test_gpu_encoder_head.py -x -q -m gpu 2>&1 | tail -12
python - <<'PY'
import sys, torch, json
sys.path.insert(0, '.')
import raft_ros_b200 as rr
import torch.nn.functional as F
torch.backends.cudnn.allow_tf32 = False
g = torch.Generator().manual_seed(1)
B, h, w = 8, 55, 128
x = torch.randn(2 * B, 128, h, w, generator=g).cuda()
W = (torch.randn(256, 128, 1, 1, generator=g) / 11).cuda(); b = torch.zeros(256).cuda()
def t(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); a, e = torch.cuda.Event(True), torch.cuda.Event(True); a.record()
    for _ in range(reps): fn()
    e.record(); torch.cuda.synchronize(); return a.elapsed_time(e) / reps * 1e3
out = {}
out["encoder_head_us"] = t(lambda: rr.encoder_head(x, W, b))
out["conv2_cudnn_fp32_us"] = t(lambda: F.conv2d(x, W, b))
torch.backends.cudnn.allow_tf32 = True
out["conv2_cudnn_tf32_us"] = t(lambda: F.conv2d(x, W, b))
torch.backends.cudnn.allow_tf32 = False
f = F.conv2d(x, W, b)
out["build_from_fp32_maps_us"] = t(lambda: rr.CorrBlock(f[:B], f[B:], precision="bf16x3"))
prep = rr.encoder_head(x, W, b)
out["build_from_prepared_us"] = t(lambda: rr.CorrBlock.from_prepared(prep, precision="bf16x3"))
out["two_step_us"] = t(lambda: rr.CorrBlock(*(lambda f: (f[:B], f[B:]))(F.conv2d(x, W, b)), precision="bf16x3"))
out["fused_head_us"] = t(lambda: rr.CorrBlock.from_prepared(rr.encoder_head(x, W, b), precision="bf16x3"))
print(json.dumps(out))
PY
