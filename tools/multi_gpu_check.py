#!/usr/bin/env python
"""Multi-GPU parity check, one rank per GPU (torchrun): blocks sharded over the ranks must give the same
codebooks and indices as one GPU, for yupvq_encode_bcn_sharded, yupvq_encode_bcn_batch_sharded (Oklab, fast-mode
sampling, SAD metric) and yupvq_encode_hdr_batch_sharded.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/multi_gpu_check.py
"""
import os
import sys

import numpy as np

ROOT = os.path.normpath(os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from yuptexture_b200 import binding, sharding, synth

    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
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

    size, k, iters = 1024, 256, 10
    texs = [synth.bc1_encode(synth.ldr_rgb(size, size, seed=200 + t)) for t in range(world + 1)]
    texs.append(synth.bc1_encode(synth.checker_rgb(size, size)))  # few distinct blocks: empty clusters, sharded reseed
    n = len(texs[0])
    first, n_local = sharding.shard_range(n, world, rank)
    cfg = binding.make_config(codebook=k, max_iterations=iters, seed=9)

    # single-GPU answers (every rank computes them on its own GPU with a second, communicator-less context)
    solo = binding.Encoder(local)
    want = []
    for t, bc in enumerate(texs):
        c = binding.make_config(codebook=k, max_iterations=iters, seed=9 + t)
        cent, idx, st = solo.encode_bcn(bc, n, 1, 3, c)
        want.append((cent, idx, st.lloyd_passes))
    solo.close()

    ok = True
    # (1) per-texture sharded call
    for t, bc in enumerate(texs):
        c = binding.make_config(codebook=k, max_iterations=iters, seed=9 + t)
        cent, idx, st = enc.encode_bcn_sharded(bc, n, first, n_local, 1, 3, c)
        ok &= np.array_equal(cent, want[t][0]) and np.array_equal(idx, want[t][1][first:first + n_local])
        ok &= st.lloyd_passes == want[t][2]
    # (2) lock-step batch
    cents, idxs, st = enc.encode_bcn_batch_sharded(texs, n, first, n_local, 1, 3, cfg)
    for t in range(len(texs)):
        ok &= np.array_equal(cents[t], want[t][0]) and np.array_equal(idxs[t], want[t][1][first:first + n_local])
    ok &= st.lloyd_passes == sum(w[2] for w in want)

    # (3) the same batch with shard-only inputs: whole textures only where this rank seeds them (t % world == rank)
    parts = [bc if t % world == rank else bc[first:first + n_local] for t, bc in enumerate(texs)]
    cents, idxs, _ = enc.encode_bcn_batch_sharded(parts, n, first, n_local, 1, 3, cfg, shard_only=[t % world != rank for t in range(len(texs))])
    for t in range(len(texs)):
        ok &= np.array_equal(cents[t], want[t][0]) and np.array_equal(idxs[t], want[t][1][first:first + n_local])
    # (4) seed == 0 (std::random_device, the reference's default) on the sharded single-texture path: rank 0's draw is
    # broadcast, so all ranks must still end with ONE codebook, and the indices must be the exact argmin against it
    c0 = binding.make_config(codebook=k, max_iterations=iters, seed=0)
    cent0, idx0, _ = enc.encode_bcn_sharded(texs[0], n, first, n_local, 1, 3, c0)
    mine = torch.from_numpy(cent0.astype(np.int32)).to(dev)
    lo, hi = mine.clone(), mine.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    ok &= bool(torch.equal(lo, hi))
    # (5) a context with a communicator still runs the rank-local entry points on its own (different texture per rank)
    c5 = binding.make_config(codebook=k, max_iterations=iters, seed=9 + rank % len(texs))
    cent5, idx5, st5 = enc.encode_bcn(texs[rank % len(texs)], n, 1, 3, c5)
    ok &= np.array_equal(cent5, want[rank % len(texs)][0]) and np.array_equal(idx5, want[rank % len(texs)][1])

    base_ok = ok
    # (6)-(8) the variants of the lock-step batch (general path): fast-mode sampling, SAD metric, both; then HDR images.
    # The checker texture has few distinct blocks: empty clusters, i.e. the host-driven sharded reseed, on every variant.
    var_ok = {}
    small = [synth.bc1_encode(synth.ldr_rgb(512, 512, seed=300 + t)) for t in range(world + 1)]
    small.append(synth.bc1_encode(synth.checker_rgb(512, 512)))
    ns = len(small[0])
    f2, nl2 = sharding.shard_range(ns, world, rank)
    solo = binding.Encoder(local)
    for name, kw in (("sampled", dict(sample_ratio=0.35)), ("sad", dict(metric=binding.METRIC_SAD)),
                     ("sad_sampled", dict(metric=binding.METRIC_SAD, sample_ratio=0.5))):
        good = True
        cfgv = binding.make_config(codebook=128, max_iterations=8, seed=40, **kw)
        cents, idxs, stv = enc.encode_bcn_batch_sharded(small, ns, f2, nl2, 1, 3, cfgv)
        passes = 0
        for t, bc in enumerate(small):
            c = binding.make_config(codebook=128, max_iterations=8, seed=40 + t, **kw)
            cent, idx, s1 = solo.encode_bcn(bc, ns, 1, 3, c)
            good &= np.array_equal(cents[t], cent) and np.array_equal(idxs[t], idx[f2:f2 + nl2])
            passes += s1.lloyd_passes
        good &= stv.lloyd_passes == passes
        var_ok[name] = bool(good)
    w, h = 512, 256
    imgs = [synth.hdr_rgb(w, h, seed=400 + t) for t in range(world + 1)]
    nh = (w // 4) * (h // 4)
    f3, nl3 = sharding.shard_range(nh, world, rank)
    for name, kw in (("hdr", dict()), ("hdr_sampled", dict(sample_ratio=0.5))):
        good = True
        cfgv = binding.make_config(codebook=600, max_iterations=8, seed=50, **kw)
        cents, idxs, _ = enc.encode_hdr_batch_sharded(imgs, w, h, 3, f3, nl3, cfgv)
        for t, img in enumerate(imgs):
            c = binding.make_config(codebook=600, max_iterations=8, seed=50 + t, **kw)
            cent, idx, _ = solo.encode_hdr(img, w, h, 3, c)
            # HDR seeding sums are doubles reduced in a fixed tree on ONE rank in both runs: bit-identical
            good &= np.array_equal(cents[t].view(np.uint32), cent.view(np.uint32)) and np.array_equal(idxs[t], idx[f3:f3 + nl3])
        var_ok[name] = bool(good)
    solo.close()
    ok = base_ok and all(var_ok.values())
    vt = torch.tensor([1 if base_ok else 0] + [1 if var_ok[k2] else 0 for k2 in sorted(var_ok)], device=dev)
    dist.all_reduce(vt, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("cases: ldr_oklab=%d " % int(vt[0]) + " ".join("%s=%d" % (k2, int(vt[1 + i])) for i, k2 in enumerate(sorted(var_ok))))

    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    enc.close()
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("multi_gpu_check ok" if int(flag) else "multi_gpu_check FAILED", f"world={world} textures={len(texs)}")
    sys.exit(0 if int(flag) else 1)


if __name__ == "__main__":
    main()
