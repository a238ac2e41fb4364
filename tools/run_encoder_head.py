"""encoder_head at batch 8 (16 images of 55x128, 128 -> 256 channels), a few calls: the command ncu captures."""
import sys
import torch
sys.path.insert(0, '.')
import raft_ros_b200 as rr
g = torch.Generator().manual_seed(1)
x = torch.randn(16, 128, 55, 128, generator=g).cuda()
W = (torch.randn(256, 128, generator=g) / 11).cuda()
b = torch.zeros(256).cuda()
for _ in range(5):
    p = rr.encoder_head(x, W, b)
torch.cuda.synchronize()
a, e = torch.cuda.Event(True), torch.cuda.Event(True)
a.record()
for _ in range(20):
    p = rr.encoder_head(x, W, b)
e.record()
torch.cuda.synchronize()
print("encoder_head us", a.elapsed_time(e) / 20 * 1e3, float(p.hi.float().mean()))
