"""Experiment: lookup on a 4x4-tiled copy of the pyramid vs the row layout (time + parity)."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, ROOT)
import raft_ros_b200 as rr
from raft_ros_b200 import _lib
from raft_ros_b200.corr import build_pyramid_raw, lookup_raw
import torch.nn.functional as F
B, D, h, w, K = 8, 256, 55, 128, 12
g = torch.Generator().manual_seed(0)
f1 = torch.randn(B, D, h, w, generator=g).cuda(); f2 = torch.randn(B, D, h, w, generator=g).cuda()
coords = (rr.coords_grid(B, h, w, "cpu")[None] + torch.cumsum(0.5 * torch.randn(K, B, 2, h, w, generator=g), 0)).cuda()
lv = build_pyramid_raw(f1, f2, 4, "bf16")
def tile(t):
    P, _, hl, wl = t.shape
    hp, wp = (hl + 3) // 4 * 4, (wl + 3) // 4 * 4
    x = F.pad(t.view(P, hl, wl), (0, wp - wl, 0, hp - hl))
    return x.view(P, hp // 4, 4, wp // 4, 4).permute(0, 1, 3, 2, 4).contiguous()
tl = [tile(t) for t in lv]
L = _lib.lib()
def tiled(c):
    out = torch.empty(B, 324, h, w, device="cuda")
    rc = L.rc_corr_lookup_tiled_fwd(_lib.ptr_array([t.data_ptr() for t in tl]), c.data_ptr(), out.data_ptr(), B, h, w, 4, 4,
                                    torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L.rc_last_error()
    return out
a = lookup_raw(lv, coords[3], 4); b = tiled(coords[3])
print("max abs diff", float((a - b).abs().max()), "equal", bool(torch.equal(a, b)))
def timeit(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / n
print("row layout  ms/lookup", timeit(lambda: [lookup_raw(lv, coords[k], 4) for k in range(K)]) / K)
print("tiled       ms/lookup", timeit(lambda: [tiled(coords[k]) for k in range(K)]) / K)
