"""Build-only timing at the BASELINE.json geometries (aligned and ragged widths)."""
import json, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from raft_ros_b200.corr import build_pyramid_raw, lookup_raw
import raft_ros_b200 as rr
def timeit(fn, warm=3, reps=20):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); a, b = torch.cuda.Event(True), torch.cuda.Event(True); a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize(); return a.elapsed_time(b) / reps
for name, B, h, w in [("sintel 55x128", 8, 55, 128), ("kitti 47x156", 8, 47, 156), ("things 50x90", 6, 50, 90)]:
    g = torch.Generator().manual_seed(0)
    f1 = torch.randn(B, 256, h, w, generator=g).cuda(); f2 = torch.randn(B, 256, h, w, generator=g).cuda()
    coords = rr.coords_grid(B, h, w, "cuda") + torch.randn(B, 2, h, w, generator=g).cuda()
    out = {"shape": name, "B": B}
    pyr_bytes = sum(B * h * w * (h >> l) * (w >> l) for l in range(4)) * 4
    for p in ("bf16x3", "bf16"):
        ms = timeit(lambda: build_pyramid_raw(f1, f2, 4, p))
        out[p + "_build_ms"] = round(ms, 4); out[p + "_build_GBs"] = round(pyr_bytes / ms / 1e6, 1)
    lv = build_pyramid_raw(f1, f2, 4, "bf16")
    ms = timeit(lambda: lookup_raw(lv, coords, 4))
    out["lookup_ms"] = round(ms, 4); out["lookup_alg_GBs"] = round(B * h * w * 2904 / ms / 1e6, 1)
    print(json.dumps(out), flush=True)
