"""Latency of the correlation path for ONE frame pair, as the ROS node runs it (ros/scripts/main.py:117-149: batch 1,
iters = 20, optionally raft-small): build + 20 lookups, eager and replayed from a CUDA graph, against the reference's
own torch-op CorrBlock on the same GPU.  At this size a kernel takes 5-10 us and the per-call host cost decides."""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import raft_ros_b200 as rr  # noqa: E402

DEV = "cuda:0"


def wall(fn, reps=30):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3


def case(name, D, radius, h, w, K=20):
    g = torch.Generator().manual_seed(7)
    f1 = torch.randn(1, D, h, w, generator=g).to(DEV)
    f2 = torch.randn(1, D, h, w, generator=g).to(DEV)
    coords = (rr.coords_grid(1, h, w, "cpu")[None] + torch.cumsum(0.5 * torch.randn(K, 1, 2, h, w, generator=g), 0)).to(DEV)
    out = {"case": name, "grid": [h, w], "D": D, "radius": radius, "lookups": K}

    def step():
        blk = rr.CorrBlock(f1, f2, radius=radius)
        o = None
        for k in range(K):
            o = blk(coords[k])
        return o

    out["eager_ms"] = wall(step)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step()
    out["cuda_graph_ms"] = wall(graph.replay)
    # per-call device time of the per-iteration pieces at this size (CUDA events around 50 back-to-back calls)
    blk = rr.CorrBlock(f1, f2, radius=radius)
    cout = 256 if radius == 4 else 96
    cin = 4 * (2 * radius + 1) ** 2
    W = (torch.randn(cout, cin, generator=g) / cin ** 0.5).to(DEV)
    bias = torch.zeros(cout, device=DEV)

    def dev_us(fn, reps=50):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            for _ in range(reps):
                fn()
        gr.replay()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(True), torch.cuda.Event(True)
        a.record()
        gr.replay()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps * 1e3

    out["lookup_us"] = dev_us(lambda: blk(coords[0]))
    out["lookup_convc1_fused_us"] = dev_us(lambda: blk.lookup_convc1(coords[0], W, bias))
    out["build_us"] = dev_us(lambda: rr.CorrBlock(f1, f2, radius=radius), reps=10)
    try:
        from oracle import build_ref_core
        core = build_ref_core.load()
    except Exception:
        core = None
    if core is not None:
        import warnings
        torch.backends.cuda.matmul.allow_tf32 = False

        def ref():
            blk = core.corr.CorrBlock(f1, f2, radius=radius)
            for k in range(K):
                blk(coords[k])

        with torch.no_grad(), warnings.catch_warnings():
            warnings.simplefilter("ignore")
            out["reference_torch_ops_ms"] = wall(ref, reps=5)
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    case("raft-things 440x1024 (Sintel pair)", 256, 4, 55, 128)
    case("raft-small 376x1248 (KITTI frame, the ROS launch file's input)", 128, 3, 47, 156)
