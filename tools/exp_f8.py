"""fp16_fp8 build: accuracy against fp64 and build time next to bf16x3 / bf16 (one-off experiment)."""
import json, sys, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_ros_b200 as rr
from raft_ros_b200.corr import build_pyramid_tiled_raw

DEV = "cuda:0"
g = torch.Generator().manual_seed(3)
out = {}
for name, mul in (("randn", 1.0), ("randn x 1e-3", 1e-3), ("randn x 300", 300.0)):
    f1 = (mul * torch.randn(1, 256, 24, 40, generator=g)).to(DEV)
    f2 = (mul * torch.randn(1, 256, 24, 40, generator=g)).to(DEV)
    want = (torch.einsum("bdi,bdj->bij", f1.double().flatten(2), f2.double().flatten(2)) / 16.0)
    for prec in ("bf16x3", "fp16_fp8", "bf16"):
        blk = rr.CorrBlock(f1, f2, num_levels=4, radius=4, precision=prec)
        got = blk.corr_pyramid[0].double().view(1, 960, 960)
        err = (got - want)
        out[f"{name} {prec}"] = {"relmax": float(err.abs().max() / want.abs().max()), "normwise": float(err.norm() / want.norm())}
print(json.dumps(out, indent=1))
B, h, w = 8, 55, 128
f1 = torch.randn(B, 256, h, w, generator=g).to(DEV); f2 = torch.randn(B, 256, h, w, generator=g).to(DEV)
for prec in ("bf16x3", "fp16_fp8", "bf16"):
    for _ in range(3):
        lv = build_pyramid_tiled_raw(f1, f2, 4, prec)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(True), torch.cuda.Event(True)
    a.record()
    for _ in range(30):
        lv = build_pyramid_tiled_raw(f1, f2, 4, prec)
    b.record(); torch.cuda.synchronize()
    print(prec, "build ms", a.elapsed_time(b) / 30)
