"""What can this box's host absorb?  N ranks (one per GPU) each stream `--mb` MB of pinned device-to-host copies
back to back; prints per-rank and aggregate GB/s, with and without binding each rank to its GPU's NUMA node.
This is the ceiling of bench.py's end-to-end arm (every lookup result is read back: 0.88 GB per step and GPU).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/probe_d2h.py
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from raft_ros_b200 import hostbind  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=876)
    ap.add_argument("--reps", type=int, default=8)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = a.mb * 1000 * 1000 // 4
    src = torch.empty(n, dtype=torch.float32, device=dev).normal_()
    out = {"n_gpus": world, "mb_per_copy": a.mb, "reps": a.reps}
    if rank == 0:
        out["topology"] = hostbind.topology()
    import contextlib
    for mode in ("unbound", "bound"):
        ctx = hostbind.near_gpu(local, rank, world) if mode == "bound" else contextlib.nullcontext()
        with ctx:
            dst = torch.empty(n, dtype=torch.float32).pin_memory()
            dst.zero_()  # first touch while bound
        if mode == "bound" and rank == 0:
            out["bind_rank0"] = ctx.info
        for direction in ("d2h", "h2d"):
            a_, b_ = (dst, src) if direction == "d2h" else (src, dst)
            a_.copy_(b_, non_blocking=True)
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            e0.record()
            for _ in range(a.reps):
                a_.copy_(b_, non_blocking=True)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / a.reps
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            gbs_rank = n * 4 / (ms * 1e-3) / 1e9
            out[f"{mode}_{direction}"] = {"slowest_rank_ms": float(t.item()), "aggregate_gbs": world * n * 4 / (float(t.item()) * 1e-3) / 1e9,
                                          "rank0_gbs": gbs_rank}
        del dst
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
