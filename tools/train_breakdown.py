"""CUDA-event timing of the raw ops behind one training step at config 4 (400x720 crop -> 50x90, batch 6)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import raft_ros_b200 as rr  # noqa: E402
from raft_ros_b200 import corr as C  # noqa: E402

DEV = "cuda:0"


def timeit(fn, warm=3, reps=20):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(True), torch.cuda.Event(True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3  # us


def main():
    B, D, h, w = 6, 256, 50, 90
    g = torch.Generator().manual_seed(1)
    f1 = torch.randn(B, D, h, w, generator=g).to(DEV)
    f2 = torch.randn(B, D, h, w, generator=g).to(DEV)
    coords = (rr.coords_grid(B, h, w, "cpu") + 2 * torch.randn(B, 2, h, w, generator=g)).to(DEV)
    gout = torch.randn(B, 324, h, w, generator=g).to(DEV)
    out = {"config": "50x90 batch 6 bf16x3"}
    out["build_rows_us"] = timeit(lambda: C.build_pyramid_raw(f1, f2, 4, "bf16x3"))
    out["build_tiled_us"] = timeit(lambda: C.build_pyramid_tiled_raw(f1, f2, 4, "bf16x3"))
    rows = C.build_pyramid_raw(f1, f2, 4, "bf16x3")
    tiled = C.build_pyramid_tiled_raw(f1, f2, 4, "bf16x3")
    out["lookup_rows_us"] = timeit(lambda: C.lookup_raw(rows, coords, 4))
    out["lookup_tiled_us"] = timeit(lambda: C.lookup_tiled_raw(tiled, coords, 4))
    grads = C._alloc_like_levels(rows)
    out["lookup_bwd_us"] = timeit(lambda: C.lookup_bwd_raw(gout, coords, grads, 4))
    out["zero_grad_pyramid_us"] = timeit(lambda: C._alloc_like_levels(rows))
    out["build_bwd_us"] = timeit(lambda: C.build_bwd_raw(grads, f1, f2, precision="bf16x3"))
    print(json.dumps(out))


if __name__ == "__main__":
    main()
