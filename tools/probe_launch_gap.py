"""How much of a lookup's in-loop time is host launch overhead?  Times N back-to-back lookups on the
GPU (events) and the host-side cost of issuing them (wall clock, no sync inside)."""
import time
import torch
import raft_ros_b200 as rr
from raft_ros_b200 import _lib
from raft_ros_b200.corr import lookup_tiled_raw, build_pyramid_raw, lookup_raw, relayout_raw, level_shapes

dev = torch.device("cuda:0")
B, D, h, w = 8, 256, 55, 128
g = torch.Generator().manual_seed(0)
f1 = torch.randn(B, D, h, w, generator=g).to(dev)
f2 = torch.randn(B, D, h, w, generator=g).to(dev)
coords = (rr.coords_grid(B, h, w, "cpu") + 2 * torch.randn(B, 2, h, w, generator=g)).to(dev)
sigma = float(__import__("os").environ.get("PROBE_SIGMA", "2"))
coords = (rr.coords_grid(B, h, w, "cpu") + sigma * torch.randn(B, 2, h, w, generator=g)).to(dev)
rows = build_pyramid_raw(f1, f2, 4, "bf16x3")


def tiled_from_rows(f1_, f2_, L_, prec):
    return [relayout_raw(x, B, h * w, hl, wl, True) for x, (hl, wl) in zip(rows, level_shapes(h, w, 4))]


print("max |tiled - rows| =", float((lookup_tiled_raw(tiled_from_rows(0, 0, 0, 0), coords, 4) - lookup_raw(rows, coords, 4)).abs().max()))
for name, bf, lf in (("tiled", tiled_from_rows, lookup_tiled_raw), ("rows", build_pyramid_raw, lookup_raw)):
    lv = bf(f1, f2, 4, "bf16x3")
    for _ in range(5):
        lf(lv, coords, 4)
    torch.cuda.synchronize()
    n = 200
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(n):
        lf(lv, coords, 4)
    e1.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    print(f"{name}: gpu {e0.elapsed_time(e1) / n * 1e3:.1f} us/launch, host issue {(t1 - t0) / n * 1e6:.1f} us/call")
    # raw C call only, preallocated output
    out = torch.empty((B, 4 * 81, h, w), device=dev)
    pa = _lib.ptr_array([t.data_ptr() for t in lv])
    fn = _lib.lib().rc_corr_lookup_tiled_fwd if name == "tiled" else _lib.lib().rc_corr_lookup_fwd
    st = torch.cuda.current_stream().cuda_stream
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(n):
        fn(pa, coords.data_ptr(), out.data_ptr(), B, h, w, 4, 4, st)
    e1.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    print(f"{name}: C ABI only: gpu {e0.elapsed_time(e1) / n * 1e3:.1f} us/launch, host issue {(t1 - t0) / n * 1e6:.1f} us/call")

# the public class at batch 1 (the ROS node's case): host cost per lookup call
f1b, f2b, cb = f1[:1].contiguous(), f2[:1].contiguous(), coords[:1].contiguous()
with torch.no_grad():
    blk = rr.CorrBlock(f1b, f2b)
    for _ in range(10):
        blk(cb)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(n):
        blk(cb)
    e1.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
print(f"CorrBlock batch 1: gpu {e0.elapsed_time(e1) / n * 1e3:.1f} us/lookup, host issue {(t1 - t0) / n * 1e6:.1f} us/call")
