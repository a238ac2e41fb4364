"""The update block's convolutions at batch 1 and 8 (55x128): rc_conv2d_fwd against cuDNN (TF32 as torch defaults it, and
fp32), with bias + activation fused on our side and as separate torch ops on the other."""
import json
import sys
import torch
import torch.nn.functional as F
sys.path.insert(0, '.')
import raft_ros_b200 as rr

def t(fn, reps=20):
    """device time per call: the calls are captured in a CUDA graph, so host-side launch cost does not count"""
    for _ in range(3): fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps): fn()
    g.replay(); torch.cuda.synchronize(); a, e = torch.cuda.Event(True), torch.cuda.Event(True); a.record()
    for _ in range(3): g.replay()
    e.record(); torch.cuda.synchronize(); return a.elapsed_time(e) / (3 * reps) * 1e3

LAYERS = [("convc2 3x3 256->192 relu", 256, 0, 192, 3, 3, "relu"), ("conv 3x3 (192+64)->128 relu", 192, 64, 128, 3, 3, "relu"),
          ("gru convz1 1x5 (128+256)->128 sigmoid", 128, 256, 128, 1, 5, "sigmoid"), ("gru convq2 5x1 (128+256)->128 tanh", 128, 256, 128, 5, 1, "tanh"),
          ("flow head conv1 3x3 128->256 relu", 128, 0, 256, 3, 3, "relu"), ("mask 1x1 256->576", 256, 0, 576, 1, 1, None)]
for B in (1, 8):
    for name, C1, C2, Cout, KH, KW, act in LAYERS:
        g = torch.Generator().manual_seed(1)
        a = torch.randn(B, C1, 55, 128, generator=g).cuda()
        b = torch.randn(B, C2, 55, 128, generator=g).cuda() if C2 else None
        w = (torch.randn(Cout, C1 + C2, KH, KW, generator=g) / 30).cuda(); bias = torch.zeros(Cout).cuda()
        fn = {None: lambda z: z, "relu": torch.relu, "sigmoid": torch.sigmoid, "tanh": torch.tanh}[act]
        def ref():
            x = a if b is None else torch.cat([a, b], dim=1)
            return fn(F.conv2d(x, w, bias, padding=(KH // 2, KW // 2)))
        torch.backends.cudnn.allow_tf32 = True
        us_tf32 = t(ref)
        torch.backends.cudnn.allow_tf32 = False
        us_fp32 = t(ref)
        us_ours = t(lambda: rr.conv2d_act(a if b is None else (a, b), w, bias, act))
        flops = 2.0 * B * 55 * 128 * Cout * (C1 + C2) * KH * KW
        print(json.dumps({"layer": name, "B": B, "ours_us": round(us_ours, 1), "cudnn_tf32_us": round(us_tf32, 1), "cudnn_fp32_us": round(us_fp32, 1),
                          "ours_tflops": round(flops / us_ours / 1e6, 1)}), flush=True)

# a GRU half (core/update.py:41-46) at 55x128: two conv2d_fused launches against the reference's torch ops
for B in (1, 8):
    g = torch.Generator().manual_seed(2)
    h = torch.tanh(torch.randn(B, 128, 55, 128, generator=g)).cuda()
    x = torch.randn(B, 256, 55, 128, generator=g).cuda()
    wz, wr, wq = ((torch.randn(128, 384, 1, 5, generator=g) / 40).cuda() for _ in range(3))
    bz = br = bq = torch.zeros(128).cuda()
    wzr, bzr = torch.cat([wz, wr]), torch.cat([bz, br])
    def ref_half():
        hx = torch.cat([h, x], dim=1)
        z = torch.sigmoid(F.conv2d(hx, wz, bz, padding=(0, 2)))
        r = torch.sigmoid(F.conv2d(hx, wr, br, padding=(0, 2)))
        q = torch.tanh(F.conv2d(torch.cat([r * h, x], dim=1), wq, bq, padding=(0, 2)))
        return (1 - z) * h + z * q
    def our_half():
        zr = rr.conv2d_fused((h, x), wzr, bzr, "sigmoid", epilogue="mul_from", split=128, aux1=h)
        return rr.conv2d_fused((zr[:, 128:], x), wq, bq, "tanh", epilogue="gru", aux1=zr[:, :128], aux2=h)
    torch.backends.cudnn.allow_tf32 = True
    us_ref = t(ref_half)
    us_ours = t(our_half)
    us_zr = t(lambda: rr.conv2d_fused((h, x), wzr, bzr, "sigmoid", epilogue="mul_from", split=128, aux1=h))
    print(json.dumps({"layer": "GRU half 1x5 (z, r, r*h, q, blend)", "B": B, "ours_us": round(us_ours, 1), "ours_zr_us": round(us_zr, 1),
                      "torch_cudnn_tf32_us": round(us_ref, 1)}), flush=True)

# convf1: the 7x7 convolution of the flow field + ReLU (core/update.py:86,92)
for B in (1, 8):
    g = torch.Generator().manual_seed(4)
    fl = torch.randn(B, 2, 55, 128, generator=g).cuda()
    w7 = (torch.randn(128, 2, 7, 7, generator=g) / 10).cuda(); b7 = torch.zeros(128).cuda()
    torch.backends.cudnn.allow_tf32 = True
    us_ref = t(lambda: F.relu(F.conv2d(fl, w7, b7, padding=3)))
    us_ours = t(lambda: rr.small_conv_relu(fl, w7, b7))
    print(json.dumps({"layer": "convf1 7x7 2->128 + ReLU (CUDA cores, fp32)", "B": B, "ours_us": round(us_ours, 1), "cudnn_tf32_us": round(us_ref, 1)}), flush=True)
