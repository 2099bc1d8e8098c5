"""Host-side logic of the multi-GPU path on CPU: index partitioning and the record all-gather over
gloo with world_size 2 (the GPU box runs the same code over NCCL)."""
from __future__ import annotations

import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from underwater_image_enhancement_b200 import shard


def test_shard_ranges_cover_everything_once():
    for total in (0, 1, 5, 256, 100_000, 100_003):
        for world in (1, 2, 3, 4, 8):
            spans = [shard.shard_range(total, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
                assert a1 == b0 and a0 <= a1
            assert max(b - a for a, b in spans) <= shard.shard_capacity(total, world)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, total, strategies, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard.shard_range(total, world, rank)
    idx = torch.arange(lo, hi, dtype=torch.float64)
    # a record that encodes its own global index, so order and content can be checked
    scores = idx[:, None, None] * 10 + torch.arange(strategies, dtype=torch.float64)[None, :, None] \
        + torch.arange(9, dtype=torch.float64)[None, None, :] / 10
    best = (torch.arange(lo, hi) % strategies).to(torch.int32)
    types = (torch.arange(lo, hi) % 3).to(torch.int32)
    rec = shard.pack_records(scores, best, types, shard.shard_capacity(total, world))
    got = shard.gather_records(rec)
    s, b, t = shard.unpack_records(got, strategies)
    # NumPy copies: tensors travel through the queue as shared-memory handles that die with the worker
    q.put((rank, s.numpy().copy(), b.numpy().copy(), t.numpy().copy()))
    dist.destroy_process_group()


@pytest.mark.parametrize("total", [7, 8])
def test_gather_records_world2(total):
    world, strategies = 2, 6
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, strategies, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    idx = torch.arange(total, dtype=torch.float64)
    want = idx[:, None, None] * 10 + torch.arange(strategies, dtype=torch.float64)[None, :, None] \
        + torch.arange(9, dtype=torch.float64)[None, None, :] / 10
    for _, s, b, t in results:
        s, b, t = torch.from_numpy(s), torch.from_numpy(b), torch.from_numpy(t)
        assert torch.equal(s, want)                       # global image order, sentinels dropped
        assert torch.equal(b, (torch.arange(total) % strategies).to(torch.int32))
        assert torch.equal(t, (torch.arange(total) % 3).to(torch.int32))
