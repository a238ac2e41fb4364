"""One update-block convolution, for ncu: run_conv_one.py B C1 C2 Cout KH KW [epilogue]."""
import sys, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raft_ros_b200 as rr
B, C1, C2, Cout, KH, KW = (int(v) for v in sys.argv[1:7])
epi = sys.argv[7] if len(sys.argv) > 7 else None
g = torch.Generator().manual_seed(1)
a = torch.randn(B, C1, 55, 128, generator=g).cuda()
b = torch.randn(B, C2, 55, 128, generator=g).cuda() if C2 else None
w = (torch.randn(Cout, C1 + C2, KH, KW, generator=g) / 30).cuda()
bias = torch.zeros(Cout).cuda()
h = torch.randn(B, Cout // 2 if epi == "mul_from" else Cout, 55, 128, generator=g).cuda()
z = torch.rand(B, Cout, 55, 128, generator=g).cuda()
for _ in range(6):
    if epi == "mul_from":
        o = rr.conv2d_fused((a, b) if b is not None else a, w, bias, "sigmoid", epilogue="mul_from", split=Cout // 2, aux1=h)
    elif epi == "gru":
        o = rr.conv2d_fused((a, b) if b is not None else a, w, bias, "tanh", epilogue="gru", aux1=z, aux2=h)
    else:
        o = rr.conv2d_fused((a, b) if b is not None else a, w, bias, "relu")
torch.cuda.synchronize()
print("ok", float(o.abs().mean()))
