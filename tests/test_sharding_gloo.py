"""CPU: the N>1 host logic (person sharding + the single keypoint gather) with two gloo
ranks, plus the partition rule itself (RSN/lib/utils/dataloader.py:88-93)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from udp_pose_b200 import sharding as S


def test_partition_rule():
    for n in (0, 1, 7, 8, 9, 1000, 1001):
        for ws in (1, 2, 3, 8):
            rows = []
            for r in range(ws):
                lo, hi = S.shard_range(n, r, ws)
                assert 0 <= lo <= hi <= n and hi - lo <= S.shard_size(n, ws)
                rows += list(range(lo, hi))
            assert rows == list(range(n))  # disjoint, ordered, complete


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, ws, port, n, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(ws))
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        full = torch.arange(n * 17 * 3, dtype=torch.float32).reshape(n, 17, 3)
        mine = S.shard(full).clone()
        lo, hi = S.shard_range(n, rank, ws)
        assert mine.shape[0] == hi - lo
        got = S.gather_keypoints(mine, n)
        q.put((rank, bool(torch.equal(got, full)), tuple(got.shape)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [8, 7, 1])
def test_gather_two_ranks(n):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    assert [r[0] for r in res] == [0, 1]
    assert all(r[1] for r in res), res
    assert all(r[2] == (n, 17, 3) for r in res)


def test_single_process_is_identity():
    x = torch.zeros((5, 17, 3))
    assert S.gather_keypoints(x, 5) is x
    assert S.shard(x, 0, 1).shape[0] == 5


def _fake_decode(cfg, out, flipped, center, scale):
    """A stand-in decoder for the host logic: keypoints that depend on every input row-wise."""
    n, K = out.shape[0], out.shape[1]
    if n == 0:  # an empty shard (more ranks than persons)
        return out.new_zeros((0, K, 3))
    f = 0 if flipped is None else flipped.reshape(n, K, -1).sum(2)
    p = torch.stack([out.reshape(n, K, -1).sum(2) + f, center[:, :1].expand(n, K) * 2, scale[:, 1:].expand(n, K) * 3], 2)
    return p.float()


def _worker_decode(rank, ws, port, n, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(ws))
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        g = torch.Generator().manual_seed(5)
        out = torch.rand((n, 17, 4, 3), generator=g)
        fl = torch.rand((n, 17, 4, 3), generator=g)
        c, s = torch.rand((n, 2), generator=g), torch.rand((n, 2), generator=g)
        got = S.decode_sharded(None, out, fl, c, s, decode_fn=_fake_decode)
        want = _fake_decode(None, out, fl, c, s)
        got2 = S.decode_sharded(None, out, None, c, s, decode_fn=_fake_decode)
        q.put((rank, bool(torch.equal(got, want)), bool(torch.equal(got2, _fake_decode(None, out, None, c, s)))))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [9, 2, 1])
def test_decode_sharded_two_ranks(n):
    """decode_sharded: every rank decodes its contiguous block, the gather restores the person order
    (odd N: the last shard is padded for the collective and the padding sliced off; N=1: rank 1 is empty)."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_decode, args=(r, 2, port, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    assert [r[0] for r in res] == [0, 1] and all(r[1] and r[2] for r in res), res
