"""Secondary measurements (not the driver's bench line): BASELINE.json configs 3-5 and the GPU
comparators of BASELINE.md section 5 (the reference's own torch ops on the B200, and its
unmodified alt_cuda_corr binary from oracle/_ref).  Prints one JSON object per measurement.

    python tools/bench_configs.py [headline] [sintel32] [kitti] [train] [altsmall] [alt4k]
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import raft_ros_b200 as rr  # noqa: E402
from oracle import build_ref  # noqa: E402  (measurement only)
from oracle import torch_port as TP  # noqa: E402

DEV = "cuda:0"


def timeit(fn, warm=3, reps=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(True), torch.cuda.Event(True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def inputs(B, D, h, w, K, seed=1234):
    g = torch.Generator().manual_seed(seed)
    f1 = torch.randn(B, D, h, w, generator=g).to(DEV)
    f2 = torch.randn(B, D, h, w, generator=g).to(DEV)
    base = rr.coords_grid(B, h, w, "cpu")
    coords = (base[None] + torch.cumsum(0.5 * torch.randn(K, B, 2, h, w, generator=g), 0)).to(DEV)
    return f1, f2, coords


def fwd_case(name, B, D, h, w, K, precisions=("bf16x3", "bf16", "bf16_vol16")):
    f1, f2, coords = inputs(B, D, h, w, K)
    out = {"case": name, "B": B, "grid": [h, w], "iters": K}

    def ours(prec):
        def run():
            blk = rr.CorrBlock(f1, f2, precision=prec)
            for k in range(K):
                blk(coords[k])
        return run

    for p in precisions:
        ms = timeit(ours(p))
        out[f"ours_{p}_ms"] = ms
        out[f"ours_{p}_pairs_per_s"] = B / (ms * 1e-3)

    torch.backends.cuda.matmul.allow_tf32 = False

    def ref():
        with torch.no_grad():
            pyr = TP.build_pyramid(f1, f2, 4)
            for k in range(K):
                TP.lookup(pyr, coords[k], 4)

    try:
        ms = timeit(ref, warm=2, reps=3)
        out["torch_ops_on_gpu_ms"] = ms
        out["torch_ops_on_gpu_pairs_per_s"] = B / (ms * 1e-3)
    except torch.cuda.OutOfMemoryError:
        out["torch_ops_on_gpu_ms"] = None
    except RuntimeError as e:  # e.g. cuDNN's grid_sample refuses the (B*h*w = 469 248)-image batch of config 3 at batch 64
        out["torch_ops_on_gpu_ms"] = None
        out["torch_ops_on_gpu_error"] = str(e).splitlines()[0][:160]
    print(json.dumps(out), flush=True)


def train_case():
    """config 4: 400x720 crop -> 50x90, batch 6, 12 lookups, forward + backward to the fmaps."""
    B, D, h, w, K = 6, 256, 50, 90, 12
    f1, f2, coords = inputs(B, D, h, w, K)
    g = torch.randn(B, 324, h, w, device=DEV)
    out = {"case": "train_400x720_b6_fwd_bwd", "B": B, "grid": [h, w], "iters": K}

    def ours(prec):
        def run():
            a, b = f1.clone().requires_grad_(True), f2.clone().requires_grad_(True)
            blk = rr.CorrBlock(a, b, precision=prec)
            loss = 0
            for k in range(K):
                loss = loss + (blk(coords[k]) * g).sum()
            loss.backward()
        return run

    for p in ("bf16x3",):
        ms = timeit(ours(p), warm=2, reps=5)
        out[f"ours_{p}_ms"] = ms
        out[f"ours_{p}_pairs_per_s"] = B / (ms * 1e-3)

    def ref():
        a, b = f1.clone().requires_grad_(True), f2.clone().requires_grad_(True)
        pyr = TP.build_pyramid(a, b, 4)
        loss = 0
        for k in range(K):
            loss = loss + (TP.lookup(pyr, coords[k], 4) * g).sum()
        loss.backward()

    ms = timeit(ref, warm=2, reps=3)
    out["torch_ops_on_gpu_ms"] = ms
    out["torch_ops_on_gpu_pairs_per_s"] = B / (ms * 1e-3)
    print(json.dumps(out), flush=True)


def alt_only(precision, H=270, W=480):
    """one precision of the on-the-fly kernel alone (ncu captures)"""
    g = torch.Generator().manual_seed(1234)
    f1 = torch.randn(1, 256, H, W, generator=g).to(DEV)
    f2 = torch.randn(1, 256, H, W, generator=g).to(DEV)
    coords = (rr.coords_grid(1, H, W, "cpu") + 2.0 * torch.randn(1, 2, H, W, generator=g)).to(DEV)
    blk = rr.AlternateCorrBlock(f1, f2, num_levels=4, radius=4, precision=precision)
    print(json.dumps({"case": f"alt4k_{precision}", "ms_per_iter": timeit(lambda: blk(coords), warm=2, reps=5)}), flush=True)


def alt_case(H=270, W=480, name="alt_2160x3840"):
    """config 5: on-the-fly correlation, one iteration = 4 levels; vs the reference's binary."""
    B, D = 1, 256
    g = torch.Generator().manual_seed(1234)
    f1 = torch.randn(B, D, H, W, generator=g).to(DEV)
    f2 = torch.randn(B, D, H, W, generator=g).to(DEV)
    coords = (rr.coords_grid(B, H, W, "cpu") + 2.0 * torch.randn(B, 2, H, W, generator=g)).to(DEV)
    blk = rr.AlternateCorrBlock(f1, f2, num_levels=4, radius=4)
    out = {"case": name, "grid": [H, W]}
    ms = timeit(lambda: blk(coords), warm=2, reps=5)
    out["ours_ms_per_iter"] = ms
    blk16 = rr.AlternateCorrBlock(f1, f2, num_levels=4, radius=4, precision="bf16")
    out["ours_bf16_ms_per_iter"] = timeit(lambda: blk16(coords), warm=2, reps=5)
    ext = build_ref.load_module()
    if ext is not None:
        f1n = f1.permute(0, 2, 3, 1).contiguous()
        f2n = [x[1].permute(0, 2, 3, 1).contiguous() for x in blk.pyramid[:4]]
        cs = [(coords.permute(0, 2, 3, 1) / 2 ** i).reshape(B, 1, H, W, 2).contiguous() for i in range(4)]

        def ref():
            for i in range(4):
                ext.forward(f1n, f2n[i], cs[i], 4)

        out["reference_binary_ms_per_iter"] = timeit(ref, warm=1, reps=3)
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    which = sys.argv[1:] or ["headline", "sintel32", "kitti", "train", "alt4k"]
    if "headline" in which:
        fwd_case("sintel_440x1024_b8_k12", 8, 256, 55, 128, 12)
    if "sintel32" in which:
        fwd_case("sintel_440x1024_b8_k32", 8, 256, 55, 128, 32)
    if "kitti" in which:
        fwd_case("kitti_376x1248_b8_k24", 8, 256, 47, 156, 24)
    if "kitti64" in which:  # config 3 as BASELINE.json states it: 64 pairs on ONE GPU (18 GB pyramid)
        fwd_case("kitti_376x1248_b64_k24", 64, 256, 47, 156, 24, precisions=("bf16x3", "bf16_vol16"))
    if "train" in which:
        train_case()
    if "altsmall" in which:
        alt_case(55, 128, "alt_440x1024")
    if "alt4k" in which:
        alt_case()
    if "alt4k_x3" in which:
        alt_only("bf16x3")
    if "alt4k_bf16" in which:
        alt_only("bf16")
