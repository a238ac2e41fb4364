#!/bin/bash
set -u
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "alternate" 2>&1 | tail -4
python - <<'PY'
import torch, sys
sys.path.insert(0, '.')
import raft_ros_b200 as rr
H, W, D = 270, 480, 256
g = torch.Generator().manual_seed(5)
a1 = torch.randn(1, D, H, W, generator=g).cuda(); a2 = torch.randn(1, D, H, W, generator=g).cuda()
c4 = (rr.coords_grid(1, H, W, "cpu") + 2.0 * torch.randn(1, 2, H, W, generator=g)).cuda()
outs = {}
for prec in ("bf16x3", "fp16_fp8", "bf16"):
    blk = rr.AlternateCorrBlock(a1, a2, num_levels=4, radius=4, precision=prec)
    for _ in range(3): o = blk(c4)
    torch.cuda.synchronize(); a, e = torch.cuda.Event(True), torch.cuda.Event(True); a.record()
    for _ in range(20): o = blk(c4)
    e.record(); torch.cuda.synchronize()
    outs[prec] = o
    print(prec, "4K ms/iter", a.elapsed_time(e) / 20)
for p in ("fp16_fp8", "bf16"):
    d = outs[p] - outs["bf16x3"]
    print(p, "vs bf16x3: relmax", float(d.abs().max() / outs["bf16x3"].abs().max()), "normwise", float(d.norm() / outs["bf16x3"].norm()))
PY
