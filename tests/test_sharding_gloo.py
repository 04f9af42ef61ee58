"""World-size-2 gloo test (CPU) of the host-side sharding logic: the shard ranges tile the block set, the
seeding owners cover the batch, and all-reduced int64 fixed-point partial sums are identical to the
single-process sums whatever the rank count -- the property that makes 1/2/4/8-GPU codebooks identical."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.normpath(os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from yuptexture_b200 import sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(3)           # same data on every rank
    n, k = 10007, 37                          # ragged: n is not a multiple of world
    feats = rng.standard_normal((n, 64)).astype(np.float32) * 3
    assign = rng.integers(0, k, n)
    first, n_local = sharding.shard_range(n, world, rank)
    # scale from the global max (all-reduce max, as the library does)
    m = torch.tensor([float(np.abs(feats[first:first + n_local]).max())])
    dist.all_reduce(m, op=dist.ReduceOp.MAX)
    shift = sharding.fixed_shift(n, float(m))
    part = sharding.partial_sums(feats[first:first + n_local], assign[first:first + n_local], k, shift)
    tot = torch.from_numpy(part.copy())
    dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    whole = sharding.partial_sums(feats, assign, k, sharding.fixed_shift(n, float(np.abs(feats).max())))
    ranges = [None] * world
    dist.all_gather_object(ranges, (first, n_local))
    q.put((rank, bool(np.array_equal(tot.numpy(), whole)), ranges))
    dist.destroy_process_group()


def test_shard_ranges_tile():
    from yuptexture_b200 import sharding
    for n in (0, 1, 7, 8, 1000, 1048576):
        for w in (1, 2, 3, 4, 8):
            cover = []
            for r in range(w):
                f, m = sharding.shard_range(n, w, r)
                cover.extend(range(f, f + m)) if n <= 1000 else cover.append((f, m))
            if n <= 1000:
                assert cover == list(range(n))
            else:
                assert sum(m for _, m in cover) == n and all(cover[i][0] + cover[i][1] == cover[i + 1][0] for i in range(w - 1))
    assert sorted({sharding.seeding_owner(t, 8) for t in range(64)}) == list(range(8))


def test_fixed_point_sums_are_rank_count_independent():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    world, port = 2, 29000 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res)
    ranges = res[0][2]
    assert ranges[0][0] == 0 and ranges[0][0] + ranges[0][1] == ranges[1][0] and ranges[1][0] + ranges[1][1] == 10007
