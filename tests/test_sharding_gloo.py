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


def _xchg_worker(rank, world, port, q, words):
    """The in-kernel exchange (csrc/tail.cu) restated on the host: pushes become all_to_all, the two phases stay."""
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from yuptexture_b200 import sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(100 + rank)
    part = rng.integers(-2 ** 50, 2 ** 50, words, dtype=np.int64)          # this rank's fixed-point partial sums
    want = torch.from_numpy(part.copy())
    dist.all_reduce(want, op=dist.ReduceOp.SUM)                           # what copy + ncclAllReduce gives
    ok = True
    for force_one_shot in (True, False):
        one_shot, slice_words = sharding.exchange_plan(words, world)
        one_shot = force_one_shot
        out = sharding.exchange_push(part, world, rank, one_shot, slice_words)
        got = [None] * world
        # (the pushes travel as objects: the slices are ragged)
        dist.all_gather_object(got, out)
        slots = [got[src][rank] for src in range(world)]                  # what landed in this rank's staging slots
        if one_shot:
            total = sharding.exchange_sum_on_read(part, slots, rank)
        else:
            mine = sharding.exchange_reduce_slice(slots)                  # phase 1: totals of this rank's slice ...
            allsl = [None] * world
            dist.all_gather_object(allsl, mine)                           # ... pushed to every rank's result buffer
            total = np.concatenate(allsl)
        ok = ok and np.array_equal(total, want.numpy())
    q.put((rank, bool(ok)))
    dist.destroy_process_group()


@pytest.mark.parametrize("words", [2 * (37 * 65 + 1), 8 * (512 * 65 + 1) + 3])
def test_exchange_step_equals_allreduce(words):
    """Both forms of the exchange step (one shot with sum-on-read, two shots over owner slices; ragged last slice) give the
    all-reduce's int64 sums bit for bit."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    world, port = 2, 31000 + (os.getpid() + words) % 2000
    procs = [ctx.Process(target=_xchg_worker, args=(r, world, port, q, words)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok in res)


def test_exchange_plan_matches_library_rule():
    from yuptexture_b200 import sharding
    assert sharding.exchange_plan(2 * (512 * 65 + 1), 2) == (True, 512 * 65 + 1)          # 2 GPUs, 2 textures: one shot
    assert sharding.exchange_plan(4 * (512 * 65 + 1), 4)[0] is True                       # 4 x 4: 3.2 MB of copies
    assert sharding.exchange_plan(8 * (512 * 65 + 1), 8)[0] is False                      # 8 x 8: 14.9 MB -> two shots
    w = 8 * (512 * 65 + 1)
    assert sharding.exchange_plan(w, 8)[1] * 8 >= w
