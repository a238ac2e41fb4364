"""CPU suite, world_size 2 over gloo: the host-side logic of the multi-GPU path (batch sharding
of frame pairs, gather for evaluation, max-over-ranks timing).  No CUDA involved."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from raft_ros_b200.shard import gather_batch, max_over_ranks, shard_batch, shard_bounds


def test_shard_bounds_partition():
    for n in [0, 1, 7, 8, 64, 65]:
        for world in [1, 2, 3, 8]:
            spans = [shard_bounds(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [e - b for b, e in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_bounds(8, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, num_pairs, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full = torch.arange(num_pairs * 6, dtype=torch.float32).view(num_pairs, 2, 3)
        mine = shard_batch(full, rank, world)
        # each rank "computes" on its own pairs only (stand-in for build + lookups)
        local = mine * 2 + 1
        whole = gather_batch(local, num_pairs)
        ok = torch.equal(whole, full * 2 + 1)
        slow = max_over_ranks(float(rank + 1), "cpu")
        q.put((rank, ok, mine.shape[0], slow))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("num_pairs", [8, 5])
def test_sharded_batch_roundtrip_world2(num_pairs):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, num_pairs, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    res.sort()
    assert all(ok for _, ok, _, _ in res)
    assert sum(n for _, _, n, _ in res) == num_pairs
    assert all(slow == float(world) for _, _, _, slow in res)  # max over ranks, seen by every rank


def test_shard_rows_partition():
    from raft_ros_b200.shard import shard_rows
    for rows in [1, 8, 43, 270, 271]:
        for world in [1, 2, 3, 8]:
            spans = [shard_rows(rows, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == rows
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            assert all(b % 16 == 0 or b == rows for b, _ in spans)  # bands start on a query-tile row
    assert [shard_rows(270, r, 8) for r in range(8)][0] == (0, 48)


def _rows_worker(rank, world, port, rows, q):
    from raft_ros_b200.shard import gather_rows, shard_rows
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full = torch.arange(2 * 3 * rows * 5, dtype=torch.float32).view(2, 3, rows, 5)
        b, e = shard_rows(rows, rank, world)
        local = full[:, :, b:e] * 3 - 1          # stand-in for AlternateCorrBlock(fmap1[:, :, b:e], fmap2)(coords[:, :, b:e])
        whole = gather_rows(local.contiguous(), rows)
        q.put((rank, torch.equal(whole, full * 3 - 1), e - b))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("rows", [43, 8])
def test_sharded_rows_roundtrip_world2(rows):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rows_worker, args=(r, world, port, rows, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res) and sum(n for _, _, n in res) == rows
