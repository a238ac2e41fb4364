"""BASELINE.json config 5 split over the GPUs of one box by QUERY ROWS: one 2160x3840
pair, feature grid 270x480, on-the-fly correlation, 12 iterations.  Every rank receives the full fmap2 (one NCCL
broadcast: the only exchange; the fmap2 pyramid is then pooled and converted locally) and its band of fmap1 rows and
of the coordinates, and computes its band of the (1, 324, 270, 480) result; nothing crosses NVLink per iteration.
Warm start (`flow_init`, core/raft.py:118-119) is just a different coords tensor: coords_grid + flow_init.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/alt_rows_sharded.py
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import raft_ros_b200 as rr  # noqa: E402
from raft_ros_b200.shard import gather_rows, max_over_ranks, shard_rows  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=12)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--precision", default="bf16x3")
    ap.add_argument("--check", action="store_true", help="gather the bands and compare with the unsharded call on rank 0")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    B, D, H, W, K = 1, 256, 270, 480, a.iters
    g = torch.Generator().manual_seed(1234)
    f1 = torch.randn(B, D, H, W, generator=g)
    f2_src = torch.randn(B, D, H, W, generator=g)
    flow_init = 3.0 * torch.randn(B, 2, H, W, generator=g)           # warm start from the previous frame
    steps = torch.cumsum(0.5 * torch.randn(K, B, 2, H, W, generator=g), 0)
    coords = rr.coords_grid(B, H, W, "cpu")[None] + flow_init[None] + steps
    b, e = shard_rows(H, rank, world)
    f1_band = f1[:, :, b:e].contiguous().to(dev)
    coords_band = coords[:, :, :, b:e].contiguous().to(dev)
    f2 = f2_src.to(dev) if rank == 0 else torch.empty(B, D, H, W, device=dev)

    def ev():
        return torch.cuda.Event(enable_timing=True)

    def timed(fn, reps):
        fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0, t1 = ev(), ev()
        t0.record()
        for _ in range(reps):
            r = fn()
        t1.record()
        torch.cuda.synchronize()
        ms = t0.elapsed_time(t1) / reps
        return (max_over_ranks(ms, dev) if world > 1 else ms), r

    bcast_ms = 0.0
    if world > 1:
        bcast_ms, _ = timed(lambda: dist.broadcast(f2, 0), 5)
    setup_ms, blk = timed(lambda: rr.AlternateCorrBlock(f1_band, f2, num_levels=4, radius=4, precision=a.precision), 3)

    def iterate():
        out = None
        for k in range(K):
            out = blk(coords_band[k]) if e > b else None
        return out

    ms, out = timed(iterate, a.reps)
    res = {"case": "config5 2160x3840 on-the-fly, query rows sharded", "n_gpus": world, "iters": K, "precision": a.precision,
           "rows_per_rank": [list(shard_rows(H, r, world)) for r in range(world)], "ms_per_iter": ms / K,
           "ms_12_iters": ms, "fmap2_broadcast_ms": bcast_ms, "block_setup_ms": setup_ms,
           "scaling": "strong (one pair, fixed work)"}
    if a.check:
        whole = gather_rows(out, H) if world > 1 else out
        if rank == 0:
            ref = rr.AlternateCorrBlock(f1.to(dev), f2, num_levels=4, radius=4, precision=a.precision)(coords[K - 1].to(dev))
            res["bit_identical_to_unsharded"] = bool(torch.equal(whole, ref))
            res["max_abs_diff"] = float((whole - ref).abs().max())
    if rank == 0:
        print(json.dumps(res), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
