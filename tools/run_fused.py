"""CorrBlock.lookup_convc1 at batch 8 @ 55x128 (the command ncu captures): python tools/run_fused.py [precision]"""
import sys
import torch
sys.path.insert(0, '.')
import raft_ros_b200 as rr
prec = sys.argv[1] if len(sys.argv) > 1 else "bf16x3"
g = torch.Generator().manual_seed(1)
B, D, h, w = 8, 256, 55, 128
f1 = torch.randn(B, D, h, w, generator=g).cuda()
f2 = torch.randn(B, D, h, w, generator=g).cuda()
W = (torch.randn(256, 324, generator=g) / 18).cuda()
b = torch.zeros(256).cuda()
coords = (rr.coords_grid(B, h, w, "cpu") + 1.5 * torch.randn(B, 2, h, w, generator=g)).cuda()
flow = torch.randn(B, 2, h, w, generator=g).cuda()
mask = torch.randn(B, 576, h, w, generator=g).cuda()
blk = rr.CorrBlock(f1, f2, precision=prec)
for _ in range(5):
    o = blk.lookup_convc1(coords, W, b)
    u = rr.upsample_flow(flow, mask)
torch.cuda.synchronize()
print("ok", float(o.mean()), float(u.mean()))
