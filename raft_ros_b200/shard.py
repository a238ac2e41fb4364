"""Batch sharding of frame pairs over the GPUs of one box.

The correlation path has no exchange step: every frame pair is independent
(reference core/corr.py:21-22,54-58 keep batch as the leading dimension and the reference
itself scatters the batch with nn.DataParallel, train.py:138).  So one process per GPU owns a
contiguous slice of the pairs, builds and looks up its own volumes, and nothing crosses
NVLink on this path.  These helpers are the host-side bookkeeping, used by bench.py and tested
on CPU with the gloo backend (tests/test_shard.py).
"""
from __future__ import annotations

from typing import List, Tuple

import torch


def shard_bounds(num_pairs: int, rank: int, world: int) -> Tuple[int, int]:
    """[begin, end) of the pairs rank `rank` owns: contiguous, sizes differ by at most one,
    the first `num_pairs % world` ranks take the extra pair."""
    if world < 1 or not (0 <= rank < world) or num_pairs < 0:
        raise ValueError(f"bad shard request: num_pairs={num_pairs}, rank={rank}, world={world}")
    base, extra = divmod(num_pairs, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def shard_batch(t: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    """The slice of a (num_pairs, ...) tensor that `rank` owns (a view, no copy)."""
    b, e = shard_bounds(t.shape[0], rank, world)
    return t[b:e]


def gather_batch(local: torch.Tensor, num_pairs: int, group=None) -> torch.Tensor:
    """Inverse of shard_batch for results that must be assembled on every rank (evaluation
    code that wants the whole batch of flow fields).  Not on the hot path: uses
    torch.distributed.all_gather on padded shards, then drops the padding."""
    import torch.distributed as dist

    world = dist.get_world_size(group)
    sizes = [shard_bounds(num_pairs, r, world) for r in range(world)]
    cap = max(e - b for b, e in sizes)
    pad = torch.zeros((cap,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    bufs: List[torch.Tensor] = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad, group=group)
    return torch.cat([bufs[r][: e - b] for r, (b, e) in enumerate(sizes)], dim=0)


def shard_rows(num_rows: int, rank: int, world: int, align: int = 16) -> Tuple[int, int]:
    """[begin, end) of the query ROWS of one large frame that `rank` owns: contiguous
    bands whose starts are multiples of `align` (the on-the-fly kernel's query tile is 16 rows -- a CTA pair -- so
    aligned bands give bit-identical results to the unsharded call), sizes as even as that allows; trailing ranks may own nothing when
    the frame has fewer tile rows than ranks.  The band of fmap1 / coords goes with the FULL fmap2:
    `AlternateCorrBlock(fmap1[:, :, b:e], fmap2)(coords[:, :, b:e])`."""
    if world < 1 or not (0 <= rank < world) or num_rows < 0 or align < 1:
        raise ValueError(f"bad shard request: num_rows={num_rows}, rank={rank}, world={world}, align={align}")
    units = (num_rows + align - 1) // align
    b, e = shard_bounds(units, rank, world)
    return min(b * align, num_rows), min(e * align, num_rows)


def gather_rows(local: torch.Tensor, num_rows: int, align: int = 16, group=None) -> torch.Tensor:
    """Assemble (B, C, rows_r, W) bands into the (B, C, num_rows, W) result on every rank (all_gather on padded bands)."""
    import torch.distributed as dist

    world = dist.get_world_size(group)
    sizes = [shard_rows(num_rows, r, world, align) for r in range(world)]
    cap = max(e - b for b, e in sizes)
    B, C, _, W = local.shape
    pad = torch.zeros((B, C, cap, W), dtype=local.dtype, device=local.device)
    pad[:, :, : local.shape[2]] = local
    bufs: List[torch.Tensor] = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad, group=group)
    return torch.cat([bufs[r][:, :, : e - b] for r, (b, e) in enumerate(sizes)], dim=2)


def max_over_ranks(value: float, device, group=None) -> float:
    """Device-side timing convention of bench.py: a step takes as long as its slowest rank."""
    import torch.distributed as dist

    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())
