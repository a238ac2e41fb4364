"""BASELINE.json config 4: FlyingThings crop 400x720, batch 6 per GPU, corr build + 12 lookups
forward AND backward, with DDP / NCCL gradient all-reduce (replaces the reference's
nn.DataParallel, train.py:138).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P tools/train_step_ddp.py [--steps 20]

The correlation operators hold no parameters, so what is all-reduced is the gradient of the
network that PRODUCES the feature maps.  The encoders are out of this repository's scope; a
stand-in with the reference model's parameter count (5 257 536 fp32) is
used: a 1x1 conv head (256->256) that the corr gradients actually flow through, plus a dormant
block that only adds all-reduce bytes.  The loss is a fixed random projection of the 12 lookup
outputs.  Reported: step time (max over ranks, CUDA events) and pairs/s for the whole job.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import raft_ros_b200 as rr  # noqa: E402
from raft_ros_b200.shard import max_over_ranks  # noqa: E402

REF_PARAMS = 5257536


class StandInEncoder(nn.Module):
    def __init__(self, dim=256):
        super().__init__()
        self.head = nn.Conv2d(dim, dim, 1)
        self.rest = nn.Parameter(torch.zeros(REF_PARAMS - sum(p.numel() for p in self.head.parameters())))

    def forward(self, x):
        return self.head(x) + 0.0 * self.rest[:1].view(1, 1, 1, 1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=6)
    ap.add_argument("--iters", type=int, default=12)
    ap.add_argument("--precision", default="bf16x3")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    B, D, h, w, K = a.batch, 256, 50, 90, a.iters
    g = torch.Generator().manual_seed(1234 + rank)
    x1 = torch.randn(B, D, h, w, generator=g).to(dev)
    x2 = torch.randn(B, D, h, w, generator=g).to(dev)
    coords = (rr.coords_grid(B, h, w, "cpu")[None] + torch.cumsum(0.5 * torch.randn(K, B, 2, h, w, generator=g), 0)).to(dev)
    proj = torch.randn(B, 324, h, w, generator=g).to(dev) / 324.0
    torch.manual_seed(1234)
    model = StandInEncoder(D).to(dev)
    if world > 1:
        model = nn.parallel.DistributedDataParallel(model, device_ids=[local])
    opt = torch.optim.AdamW(model.parameters(), lr=1e-5)

    def step(sync=True):
        opt.zero_grad(set_to_none=True)
        net = model if (sync or world == 1) else model.module  # the bare module: same step, no gradient exchange
        f1, f2 = net(x1).float(), net(x2).float()
        blk = rr.CorrBlock(f1, f2, radius=4, precision=a.precision)
        loss = 0
        for k in range(K):
            loss = loss + (blk(coords[k]) * proj).sum()
        loss.backward()
        opt.step()
        return loss

    def timed(fn, n):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(n):
            r = fn()
        e1.record()
        torch.cuda.synchronize()
        ms_ = e0.elapsed_time(e1) / n
        return (max_over_ranks(ms_, dev) if world > 1 else ms_), r

    for _ in range(a.warmup):
        step()
    ms, loss = timed(step, a.steps)
    out = {"case": "config4 train step (corr fwd+bwd, DDP NCCL all-reduce of 5.26M params)",
           "n_gpus": world, "batch_per_gpu": B, "iters": K, "precision": a.precision,
           "ms_per_step": ms, "pairs_per_s": world * B / (ms * 1e-3), "loss": float(loss.detach())}
    if world > 1:
        # where the all-reduce goes: the same step without gradient synchronisation, and the all-reduce alone
        step(False)
        ms_nosync, _ = timed(lambda: step(False), a.steps)
        flat = torch.zeros(REF_PARAMS, device=dev)
        dist.all_reduce(flat)
        ms_ar, _ = timed(lambda: dist.all_reduce(flat), 20)
        out.update(ms_per_step_no_sync=ms_nosync, allreduce_alone_ms=ms_ar, allreduce_bytes=REF_PARAMS * 4,
                   allreduce_exposed_ms=ms - ms_nosync,
                   note="DDP reduces its single 21 MB bucket when the last gradient (the 1x1 head, produced by rc_corr_build_bwd "
                        "at the very end of the backward) is ready: there is no later backward work to hide it behind, so "
                        "exposed ~= alone; it is small against the step")
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
