#!/usr/bin/env python
"""The in-kernel peer exchange must not hang when the ranks fall out of step (tail.cu).

Run under torchrun on >= 2 GPUs with YUPVQ_XCHG_DESYNC_RANK=1 and a short YUPVQ_XCHG_TIMEOUT_MS: rank 1 skips one exchange
number at pass 2, so its wait can never complete, and from the next pass on it no longer signals, so every other rank's
wait cannot complete either.  Every rank must come back from the call with the time-out error (not hang, not return
garbage as a success), and the next call on the same contexts -- now on the copy + ncclAllReduce path -- must give
the single-GPU answer.

    YUPVQ_XCHG_DESYNC_RANK=1 YUPVQ_XCHG_TIMEOUT_MS=300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 \
        --master-addr 127.0.0.1 tools/xchg_timeout_check.py
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.normpath(os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from yuptexture_b200 import binding, sharding, synth

    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    assert world >= 2 and os.environ.get("YUPVQ_XCHG_DESYNC_RANK") == "1"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    enc = binding.Encoder(local)
    uid = torch.zeros(128, dtype=torch.uint8)
    if rank == 0:
        uid = torch.from_numpy(enc.comm_unique_id().copy())
    uid = uid.to(dev)
    dist.broadcast(uid, 0)
    enc.comm_init(uid.cpu().numpy(), world, rank)

    bc = synth.bc1_encode(synth.ldr_rgb(1024, 1024, seed=321))
    n, k = len(bc), 256
    first, n_local = sharding.shard_range(n, world, rank)
    cfg = binding.make_config(codebook=k, max_iterations=10, seed=5)
    solo = binding.Encoder(local)
    want_cent, want_idx, _ = solo.encode_bcn(bc, n, 1, 3, cfg)
    solo.close()

    t0 = time.time()
    msg = ""
    try:
        enc.encode_bcn_sharded(bc, n, first, n_local, 1, 3, cfg)
    except RuntimeError as e:
        msg = str(e)
    took = time.time() - t0
    timed_out = "timed out" in msg
    # the contexts stay usable: the peer path is off now, the NCCL path gives the single-GPU answer
    cent, idx, _ = enc.encode_bcn_sharded(bc, n, first, n_local, 1, 3, cfg)
    same = np.array_equal(cent, want_cent) and np.array_equal(idx, want_idx[first:first + n_local])
    same = same and enc.exchange_path() == "nccl"
    flag = torch.tensor([int(timed_out and same and took < 30.0)], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    print(f"rank {rank}: error={msg!r} after {took:.2f} s; next call identical to single GPU: {same}", flush=True)
    enc.close()
    dist.destroy_process_group()
    if rank == 0:
        print("xchg_timeout_check ok" if int(flag.item()) == 1 else "xchg_timeout_check FAILED", flush=True)
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
