import os, sys, torch
import torch.nn.functional as F
sys.path.insert(0, '.')
import raft_ros_b200 as rr
def run(B, Cin, Cout, H, W, KH, KW):
    g = torch.Generator().manual_seed(1)
    x = torch.randn(B, Cin, H, W, generator=g).cuda()
    w = (torch.randn(Cout, Cin, KH, KW, generator=g) / (Cin*KH*KW)**0.5).cuda()
    want = F.conv2d(x.double(), w.double(), padding=(KH//2, KW//2)).float()
    got = rr.conv2d_act(x, w, None, None)
    err = float((got - want).abs().max() / want.abs().max())
    print(f"dbg={os.environ.get('RAFT_CONV_DBG','0')} {KH}x{KW} Cin={Cin} Cout={Cout} {H}x{W}: relmax {err:.3e}")
run(1, 64, 64, 4, 128, 1, 1)
run(1, 64, 64, 4, 128, 1, 3)
run(1, 64, 64, 4, 128, 3, 1)
run(1, 64, 64, 4, 128, 3, 3)
run(1, 256, 192, 8, 128, 3, 3)
run(1, 64, 64, 4, 100, 1, 5)
