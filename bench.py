#!/usr/bin/env python
"""bench.py -- VQ blocks/sec (seeding + Lloyd + assign) on 4096x4096 BC1, the BASELINE.json metric.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one pass of the whole VQ hot path over one synthetic 4096^2 BC1 texture (1,048,576
blocks, K = 512, 32 Lloyd iterations max, Oklab metric): BCn decode -> Oklab features, K-Means++
seeding, Lloyd iterations, centroid conversion, final assignment.  N > 1 (torchrun, one rank per
GPU): weak scaling -- N textures per step in lock step (BASELINE config 5's shape), every texture's blocks
sharded over the N GPUs, seeding texture-parallel (texture t on rank t % N), and ONE exchange per Lloyd pass
carrying the K x 65 fixed-point sums/counts of all N textures (the exchange step north_star names): inside the pass-tail
kernel over NVLink peer memory, or one NCCL all-reduce where peers cannot be mapped (YUPVQ_NO_PEER_XCHG=1 forces it).
--textures T / --placement texture: BASELINE config 5 as stated (64 textures) and the zero-collective placement.

`value`  : blocks/s with the BCn bytes already resident in HBM and outputs left in HBM.
`e2e`    : the same through the C ABI with HOST (pinned) buffers: H2D of the BCn bytes and D2H of the
           centroids + indices inside the timed region.
--impl reference : the unmodified reference (oracle/_ref) on all host cores, one full pass over the same 4096^2
           texture per timed step (8.6 s on 16 cores); the scalar CPU port on a crop where oracle/_ref is not built.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

TEX = 4096          # BASELINE config 2
QUALITY = 0.8       # -> K = 512
MAX_ITER = 32
SEED = 42
CPU_SAMPLE = 4096   # our arm's cpu_baseline: the reference on the whole bench texture (one pass, ~9 s on the box's 16 cores)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "src": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "src": "fallback"}


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self._stop, self._t = index, [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                parts = [x.strip() for x in out.strip().split(",")]
                if len(parts) >= 6:
                    self.rows.append(parts)
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


def make_texture(seed):
    from yuptexture_b200 import synth
    return synth.bc1_encode(synth.ldr_rgb(TEX, TEX, seed=seed))


def cpu_reference_step(px_sample, threads):
    """One timed pass of the reference's own CPU path (BuildCodebook + QuantizeBlocks) on the sample."""
    from oracle import refbind, orcbind
    k = 512
    if refbind.available():
        t0 = time.perf_counter()
        cent, _, _ = refbind.build_codebook(px_sample, 3, 1, QUALITY, MAX_ITER, SEED, threads=threads)
        idx = refbind.quantize_blocks(px_sample, 3, cent, threads=threads)
        idx = idx[0] if isinstance(idx, tuple) else idx
        return time.perf_counter() - t0, "reference", cent, idx
    t0 = time.perf_counter()
    cent, _ = orcbind.build_codebook(px_sample, 3, k, MAX_ITER, SEED, threads=1)
    idx = orcbind.quantize_blocks(px_sample, 3, cent)
    return time.perf_counter() - t0, "port", cent, idx


def parity_report(enc, px, cent_ref, idx_ref):
    """SURVEY 8d parity gates on the CPU sample, GPU path vs the CPU reference output of the same sample and seed."""
    from oracle import orcbind
    from yuptexture_b200 import binding, synth
    k = len(cent_ref)
    cfg = binding.make_config(codebook=k, max_iterations=MAX_ITER, seed=SEED)
    cent, st = enc.build_codebook(px, 3, cfg)
    idx, _ = enc.quantize_blocks(px, 3, cent)
    idx_on_ref, _ = enc.quantize_blocks(px, 3, cent_ref)            # indices given the reference's centroids
    cfg0 = binding.make_config(codebook=k, max_iterations=0, seed=SEED)
    seeds_gpu, _ = enc.build_codebook(px, 3, cfg0)
    seeds_cpu, _ = orcbind.build_codebook(px, 3, k, 0, SEED)         # the pinned restatement (seeds only: cheap)
    return {"sample_blocks": int(len(px)), "index_agreement_given_reference_centroids": float((idx_on_ref == idx_ref).mean()),
            "seed_list_identical": bool(np.array_equal(seeds_gpu, seeds_cpu)),
            "psnr_db_gpu": float(synth.psnr_u8(cent[idx], px)), "psnr_db_reference": float(synth.psnr_u8(cent_ref[idx_ref], px)),
            "lloyd_passes_gpu": int(st.lloyd_passes)}


def cpu_sample_pixels(bc, size=None):
    """Decoded pixel blocks of the top-left size^2 crop of the bench texture."""
    from oracle import orcbind
    size = size or CPU_SAMPLE
    bx = TEX // 4
    rows = np.arange(size // 4)[:, None] * bx + np.arange(size // 4)[None, :]
    return orcbind.decode_bcn(bc[rows.reshape(-1)], 1, 3)


def variant_parity(enc, solo, world, rank):
    """The sharded variants next to the headline path (include/yupvq.h: fast-mode sampling, SAD metric, HDR images over the
    ranks), each compared bit for bit with the communicator-less single-GPU call -- so that every N > 1 bench line, the driver's
    included, carries their multi-GPU parity and not only the builder's tools/multi_gpu_check.py run.  Collective: every rank
    calls it."""
    from yuptexture_b200 import binding, sharding, synth
    res = {}
    small = [synth.bc1_encode(synth.ldr_rgb(512, 512, seed=300 + t)) for t in range(world + 1)]
    small.append(synth.bc1_encode(synth.checker_rgb(512, 512)))     # few distinct blocks: empty clusters, sharded reseed
    ns = len(small[0])
    f2, nl2 = sharding.shard_range(ns, world, rank)
    for name, kw in (("sampled", dict(sample_ratio=0.35)), ("sad", dict(metric=binding.METRIC_SAD)),
                     ("sad_sampled", dict(metric=binding.METRIC_SAD, sample_ratio=0.5))):
        good = True
        cents, idxs, stv = enc.encode_bcn_batch_sharded(small, ns, f2, nl2, 1, 3, binding.make_config(codebook=128, max_iterations=8, seed=40, **kw))
        passes = 0
        for t, bc in enumerate(small):
            cent, idx, s1 = solo.encode_bcn(bc, ns, 1, 3, binding.make_config(codebook=128, max_iterations=8, seed=40 + t, **kw))
            good = good and np.array_equal(cents[t], cent) and np.array_equal(idxs[t], idx[f2:f2 + nl2])
            passes += s1.lloyd_passes
        res[name] = bool(good and stv.lloyd_passes == passes)
    w, h = 512, 256
    imgs = [synth.hdr_rgb(w, h, seed=400 + t) for t in range(world + 1)]
    nh = (w // 4) * (h // 4)
    f3, nl3 = sharding.shard_range(nh, world, rank)
    for name, kw in (("hdr", dict()), ("hdr_sampled", dict(sample_ratio=0.5))):
        good = True
        cents, idxs, _ = enc.encode_hdr_batch_sharded(imgs, w, h, 3, f3, nl3, binding.make_config(codebook=600, max_iterations=8, seed=50, **kw))
        for t, img in enumerate(imgs):
            cent, idx, _ = solo.encode_hdr(img, w, h, 3, binding.make_config(codebook=600, max_iterations=8, seed=50 + t, **kw))
            good = good and np.array_equal(cents[t].view(np.uint32), cent.view(np.uint32)) and np.array_equal(idxs[t], idx[f3:f3 + nl3])
        res[name] = bool(good)
    return res


def workload_config(n_tex):
    """The `config` object both arms print (identical keys and values for the same N, so the driver's same_config holds)."""
    return {"workload": f"BC1 VQ compress of {n_tex} synthetic 4096x4096 texture(s) (BASELINE config 2 per GPU): "
                        "seeding+Lloyd+assign, K=512, maxIterations=32, Oklab metric",
            "blocks_per_texture": (TEX // 4) ** 2, "textures": n_tex, "codebook": 512, "max_iterations": MAX_ITER, "seed": SEED}


def run_reference(args, rank, world):
    """The reference's own CPU implementation of the path (oracle/_ref: the unmodified sources; else the CPU port) on all
    host cores, on the GPU arm's exact workload: every timed step is one full pass over the whole 4096^2 texture
    (BuildCodebook + QuantizeBlocks, src/vq_encoder.cpp:260 / :447; ~9 s on the box's 16 cores).  The first warm-up is a
    full pass too; further warm-ups run on a 1024^2 crop.  The scalar port (no oracle/_ref) is bounded to the crop."""
    if rank != 0:
        return
    from oracle import orcbind, refbind
    bc = make_texture(1000)
    threads = os.cpu_count() or 1
    have_ref = refbind.available()
    px_crop = cpu_sample_pixels(bc, 1024)
    px = orcbind.decode_bcn(bc, 1, 3) if have_ref else px_crop
    for w in range(args.warmup):
        cpu_reference_step(px if w == 0 else px_crop, threads)
    per, builds, quants = [], [], []
    kind = "port"
    cent = None
    for _ in range(args.steps):
        if have_ref:
            t0 = time.perf_counter()
            cent, _, t_build = refbind.build_codebook(px, 3, 1, QUALITY, MAX_ITER, SEED, threads=threads)
            _, t_quant = refbind.quantize_blocks(px, 3, cent, threads=threads)
            per.append(time.perf_counter() - t0)
            builds.append(t_build)
            quants.append(t_quant)
            kind = "reference"
        else:
            per.append(cpu_reference_step(px, threads)[0])
    dt = float(np.mean(per)) if per else 0.0
    value = len(px) / dt if dt else 0.0
    breakdown = None
    if have_ref and cent is not None:
        # the K centroid encodes (A16) that BuildCodebook's time includes, timed on their own (src/vq_encoder.cpp:432-440)
        t1 = time.perf_counter()
        refbind.encode_blocks(cent, 3, 1, 1.0, threads)
        t_a16 = time.perf_counter() - t1
        # the whole-texture Compressonator BC1 pre-pass of VQBCnCompressor::Compress (src/vq_bcn_compressor.cpp:94)
        from yuptexture_b200 import synth
        img = synth.ldr_rgb(TEX, TEX, seed=1000)
        t2 = time.perf_counter()
        refbind.bcn_encode(img, 1, quality=1.0, threads=threads)
        t_pre = time.perf_counter() - t2
        tb, tq = float(np.mean(builds)), float(np.mean(quants))
        breakdown = {"bcn_encode_s": t_pre, "build_codebook_s": tb, "of_which_centroid_encodes_s": t_a16,
                     "quantize_blocks_s": tq, "vq_core_s": tb + tq - t_a16,
                     "note": "reference stages of VQBCnCompressor::Compress on the 4096^2 texture (example/example.cpp:293-304 prints "
                             "the sum); the per-block decode loop (:142-146) and zstd are not timed here"}
    sample = (f"{args.steps} full passes over the 4096^2 texture ({len(px)} blocks each, K=512, 32 iterations max), mean {dt:.2f} s"
              if have_ref else f"scalar CPU port on a 1024x1024 crop ({len(px)} blocks), mean {dt:.2f} s")
    print(json.dumps({
        "impl": "reference", "metric": "vq_blocks_per_sec", "value": value, "unit": "blocks/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(max(1, args.gpus)),
        "note": "one texture per step on the host cores whatever N is: the reference has no multi-texture parallelism, N textures cost N "
                "times this (blocks/s is unchanged)",
        "compress_breakdown": breakdown,
        "cpu_baseline": {"value": value, "unit": "blocks/s", "cores": threads if kind == "reference" else 1,
                         "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "blocks/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from yuptexture_b200 import binding, sharding

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    enc = binding.Encoder(local_rank)
    k = binding.codebook_size(QUALITY)
    n = (TEX // 4) ** 2
    n_tex = args.textures or world       # default = weak scaling: one texture's worth of blocks per GPU
    texture_parallel = args.placement == "texture" and world > 1
    cfg = binding.make_config(quality=QUALITY, max_iterations=MAX_ITER, seed=SEED)

    if world > 1 and not texture_parallel:
        uid = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            uid = torch.from_numpy(enc.comm_unique_id().copy())
        uid = uid.to(dev)
        dist.broadcast(uid, 0)
        enc.comm_init(uid.cpu().numpy(), world, rank)

    first, n_local = sharding.shard_range(n, world, rank)
    if texture_parallel:
        first, n_local = 0, n
    # which textures this rank holds: block-sharded = the whole textures it seeds + its block shard of all the others
    # (yupvq_encode_bcn_batch_sharded, include/yupvq.h); texture-parallel = whole textures t % world == rank only
    mine = [t for t in range(n_tex) if t % world == rank]
    held = mine if texture_parallel else list(range(n_tex))

    # synthetic inputs (fixed seeds 1000 + t); texture t is generated on rank t % world and its bytes (or this rank's shard of
    # them) are broadcast
    pinned_in, dev_in, whole, check_full = {}, {}, {}, {}
    tex0_host = None
    for t in range(n_tex):
        owner = t % world
        if world == 1:
            full = make_texture(1000 + t)
        else:
            buf = torch.empty(n * 8, dtype=torch.uint8, device=dev)
            if owner == rank:
                buf.copy_(torch.from_numpy(make_texture(1000 + t).reshape(-1).copy()))
            if not texture_parallel or t in (0, n_tex - 1):
                dist.broadcast(buf, owner)
            full = buf.cpu().numpy().reshape(n, 8) if (owner == rank or not texture_parallel or t in (0, n_tex - 1)) else None
        if t == 0:
            tex0_host = full
        if t in (0, n_tex - 1) and full is not None:
            check_full[t] = full
        if t not in held:
            continue
        is_whole = texture_parallel or world == 1 or owner == rank
        part = full if is_whole else full[first:first + n_local]
        whole[t] = is_whole
        pinned_in[t] = torch.from_numpy(np.ascontiguousarray(part).reshape(-1).copy()).pin_memory()
        dev_in[t] = pinned_in[t].to(dev)
    dev_cent = {t: torch.empty(k * 48, dtype=torch.uint8, device=dev) for t in held}
    dev_idx = {t: torch.empty(max(1, n_local), dtype=torch.int32, device=dev) for t in held}
    host_cent = {t: torch.empty(k * 48, dtype=torch.uint8).pin_memory() for t in held}
    host_idx = {t: torch.empty(max(1, n_local), dtype=torch.int32).pin_memory() for t in held}

    def step(resident):
        src = (lambda t: dev_in[t].data_ptr()) if resident else (lambda t: pinned_in[t].numpy())
        oc = (lambda t: dev_cent[t].data_ptr()) if resident else (lambda t: host_cent[t].numpy())
        oi = (lambda t: dev_idx[t].data_ptr()) if resident else (lambda t: host_idx[t].numpy().view(np.uint32))
        if world == 1 and n_tex == 1:
            _, _, st = enc.encode_bcn(src(0), n, 1, 3, cfg, oc(0), oi(0))
            return [st]
        if texture_parallel:   # independent textures, zero collectives: texture t runs whole on rank t % world with seed + t
            sts = []
            for t in mine:
                c = binding.make_config(quality=QUALITY, max_iterations=MAX_ITER, seed=SEED + t)
                sts.append(enc.encode_bcn(src(t), n, 1, 3, c, oc(t), oi(t))[2])
            return sts
        # lock-step batch: every texture's blocks sharded over the GPUs; one all-reduce per pass
        _, _, st = enc.encode_bcn_batch_sharded([src(t) for t in held], n, first, n_local, 1, 3, cfg, [oc(t) for t in held],
                                                [oi(t) for t in held], shard_only=[not whole[t] for t in held])
        return [st]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(resident, steps):
        launches, last = 0, None
        barrier()
        enc.timer_start()
        t0 = time.perf_counter()
        for _ in range(steps):
            for st in step(resident):
                launches += st.gpu_launches
                last = st
        ms_dev = enc.timer_stop()
        barrier()
        wall = (time.perf_counter() - t0) * 1e3
        ms = max(ms_dev, 0.0)
        if world > 1:
            tt = torch.tensor([ms, wall], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms, wall = float(tt[0]), float(tt[1])
        return ms, wall, launches, last

    for _ in range(max(3, args.warmup)):
        step(True)
    with ClockSampler(local_rank) as clocks:
        ms_dev, wall_dev, launches, last = timed(True, args.steps)
    for _ in range(2):
        step(False)
    ms_e2e, wall_e2e, _, last_e2e = timed(False, args.steps)
    exchange_path = enc.exchange_path()   # of the last timed step, before any checking call below

    total_blocks = n * n_tex
    value = total_blocks * args.steps / (ms_dev * 1e-3)
    e2e = total_blocks * args.steps / (ms_e2e * 1e-3)

    # one more real step (all ranks: it contains collectives) with the front-end launches bracketed by events
    enc.profile(True)
    step(True)
    ms_fe, n_fe = enc.profile_read()
    stage_ms, stage_passes = enc.profile_stages()
    enc.profile(False)

    # ---- multi-GPU parity, carried by every N > 1 line: this rank's outputs of the first and the last texture it holds are
    # compared bit for bit with a communicator-less single-GPU yupvq_encode_bcn of the same texture and seed
    multi = None
    if world > 1:
        solo = binding.Encoder(local_rank)
        ok, checked = True, 0
        for t in sorted({0, n_tex - 1}):
            if t not in held or t not in check_full:
                continue
            c = binding.make_config(quality=QUALITY, max_iterations=MAX_ITER, seed=SEED + t)
            cent_s, idx_s, _ = solo.encode_bcn(check_full[t], n, 1, 3, c)
            ok = ok and np.array_equal(dev_cent[t].cpu().numpy().reshape(k, 48), cent_s)
            ok = ok and np.array_equal(dev_idx[t].cpu().numpy().view(np.uint32)[:n_local], idx_s[first:first + n_local])
            checked += 1
        variants = None if (args.no_extras or texture_parallel) else variant_parity(enc, solo, world, rank)
        solo.close()
        tt = torch.tensor([1.0 if ok else 0.0, float(checked), 1.0 if checked else 0.0], dtype=torch.float64, device=dev)
        mn = tt.clone()
        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
        dist.all_reduce(tt, op=dist.ReduceOp.SUM)
        multi = {"multi_gpu_identical": bool(mn[0] == 1.0), "ranks_checked": int(tt[2]), "comparisons": int(tt[1]),
                 "what": "centroids and this rank's index shard of textures 0 and T-1 vs a single-GPU yupvq_encode_bcn of the "
                         "same texture and seed, bit for bit, on every rank that holds them"}
        if variants is not None:
            names = sorted(variants)
            vt = torch.tensor([1.0 if variants[v] else 0.0 for v in names], dtype=torch.float64, device=dev)
            dist.all_reduce(vt, op=dist.ReduceOp.MIN)      # a variant counts only if every rank found it identical
            multi["variants_identical"] = {v: bool(vt[i] == 1.0) for i, v in enumerate(names)}
            multi["variants_what"] = ("the other sharded entry points on small inputs (512^2 BC1 x (N+2) textures incl. a checker texture "
                                      "with empty clusters; 512x256 float images), each vs its single-GPU call with seed + t, bit for bit")

    out = None
    if rank == 0:
        pk = peaks()
        # dominant kernel: the tcgen05 assignment front-end (assign_tc_kernel), every launch of it inside one more real step
        # bracketed by CUDA events on the library's stream (yupvq_profile)
        ms_a = ms_fe / max(1, n_fe)
        if world > 1 and not texture_parallel:
            blocks_per_launch = n_local * n_tex          # the lock-step batch launches the front-end once for all textures
        else:
            blocks_per_launch = n
        flop_alg = 2.0 * blocks_per_launch * k * 64        # 2 K D per block per pass
        # executed MMA work: K = 16 steps per 128 x 256 tile-chunk actually issued for this layout -- 10 for textures without
        # alpha (48 colour columns + folded norm: 3 + 3 + 4 steps), 13 with alpha (assign_tc.cu, MMA issuer)
        ksteps = 10
        flop_exec = 2.0 * blocks_per_launch * k * 16 * ksteps
        sec = ms_a * 1e-3
        ach_alg = flop_alg / sec / 1e12 if sec > 0 else 0.0
        ach_exec = flop_exec / sec / 1e12 if sec > 0 else 0.0
        # DRAM bytes per launch come from an ncu capture, which cannot be taken inside a timed run.  The committed figure is
        # used only while it describes THIS kernel: the capture file names the sha256 of csrc/assign_tc.cu it was taken from,
        # and a capture of any other source (or of another config) leaves traffic null rather than quoting a stale number.
        traffic, traffic_source = None, None
        import glob
        import hashlib
        src_path = os.path.join(ROOT, "yuptexture_b200", "csrc", "assign_tc.cu")
        src_sha = hashlib.sha256(open(src_path, "rb").read()).hexdigest()[:16] if os.path.exists(src_path) else None
        for tpath in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_traffic.json")), reverse=True):
            tj = json.load(open(tpath))
            if world == 1 and n_tex == 1 and src_sha and tj.get("kernel_source_sha16") == src_sha:
                traffic = tj["assign_tc_kernel"]["dram_bytes_per_launch"]
                traffic_source = os.path.relpath(tpath, ROOT)
                break
        peak = pk["bf16_tflops_sustained"]
        roofline = {"bound": "tensor", "kernel": "assign_tc_kernel (N x K x 64 contraction as bf16 partial products + argmax epilogue)",
                    "achieved": ach_alg, "peak": peak, "unit": "TFLOP/s", "frac": ach_alg / peak, "traffic": traffic, "traffic_source": traffic_source,
                    "frac_algorithmic": ach_alg / peak, "achieved_executed": ach_exec, "frac_executed": ach_exec / peak,
                    "executed_over_algorithmic": 16.0 * ksteps / 64.0,
                    "peak_source": pk["src"] + " bf16 sustained (kernel timed inside a step)",
                    "ms_per_launch": ms_a, "launches_timed": n_fe, "flop_per_launch_algorithmic": flop_alg,
                    "flop_per_launch_executed": flop_exec, "mma_ksteps_per_tile_chunk": ksteps,
                    "share_of_step": ms_fe / (ms_dev / args.steps) if ms_dev > 0 else None}
        phases = ("h2d", "decode", "seed", "lloyd", "finalize", "quantize", "d2h")
        bytes_in = sum(int(pinned_in[t].numel()) for t in held)
        bytes_out = (k * 48 + n_local * 4) * len(held)
        cfgd = workload_config(n_tex)
        par = "1 GPU" if world == 1 and n_tex == 1 else (
            f"{n_tex} textures, texture-parallel placement: whole textures on rank t % {world}, zero collectives" if texture_parallel else
            f"{n_tex} textures in lock step, blocks sharded over {world} GPU(s), seeding texture-parallel, one exchange of "
            f"{n_tex} x K x 65 int64 per pass (inside the pass-tail kernel over NVLink peer memory; ncclAllReduce where peers cannot be "
            f"mapped or the batch exceeds the fused path); a rank receives whole only the textures it seeds and its block shard of the others")
        out = {
            "metric": "vq_blocks_per_sec", "value": value, "unit": "blocks/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
            "scaling": "weak" if not args.textures else "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": cfgd,
            "detail": {"parallelism": par,
                       # what the library says it did on rank 0's last timed step (yupvq_exchange_path), not a static claim
                       "exchange_path": exchange_path,
                       "l2": "no explicit flush: each step streams a 330 MB working set (features + pixel planes) > 126 MB L2",
                       "lloyd_passes": last.lloyd_passes, "ms_per_texture": ms_dev / args.steps / n_tex,
                       "phases_ms": {p: getattr(last, "ms_" + p) for p in phases},
                       "wall_ms_per_step": wall_dev / args.steps,
                       "lloyd_stage_ms": {"passes": stage_passes, **stage_ms,
                                          "note": "rank 0, one profiled step: CUDA events between the four launches of every pass"}},
            "e2e": {"value": e2e, "unit": "blocks/s", "h2d_bytes_per_step": bytes_in, "d2h_bytes_per_step": bytes_out,
                    "ms_per_step": ms_e2e / args.steps, "wall_ms_per_step": wall_e2e / args.steps,
                    "note": "rank 0's bytes; pinned host buffers, copies inside the timed region"},
            "gpu_launches": launches,
            "roofline": roofline,
            "clocks": clocks.summary(),
        }
        if multi is not None:
            out["parity_multi_gpu"] = multi
        if not args.no_extras:
            # bounded CPU sample of the same workload on this box's host cores + the SURVEY 8d parity gates on it
            px = cpu_sample_pixels(tex0_host)
            threads = os.cpu_count() or 1
            dt, kind, cent_ref, idx_ref = cpu_reference_step(px, threads)
            solo = binding.Encoder(local_rank)   # no communicator: the parity sample is a single-GPU call on rank 0 only
            out["parity"] = parity_report(solo, px, cent_ref, idx_ref)
            out["cpu_baseline"] = {"value": len(px) / dt, "unit": "blocks/s", "cores": threads if kind == "reference" else 1, "kind": kind,
                                   "sample": f"the whole {CPU_SAMPLE}x{CPU_SAMPLE} bench texture ({len(px)} blocks), one pass, "
                                             f"BuildCodebook+QuantizeBlocks {dt:.2f} s"}
            if world == 1:
                out["decompress_gather"] = gather_report(solo, torch, dev, pk)
                out["bcn_encode_gpu"] = bcn_encode_report(solo, torch, dev, pk)
                out["compress_breakdown"] = compress_report()
            solo.close()
        print(json.dumps(out))
    enc.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out


def gather_report(enc, torch, dev, pk):
    """the DecompressToBCn gather bc[i] = codebook[idx[i]] (src/vq_bcn_compressor.cpp:350-358; README.md:18 quotes
    7-11 ms for 4096x4096 on the CPU).  Device-resident: CUDA events around the C-ABI call; host: the same call with pinned
    host buffers (4 MB of indices in, 8-16 MB of blocks out over PCIe)."""
    rep = {"reference_readme_ms_4096": "7-11 (CPU memcpy loop, README.md:18)", "peak_gbs": pk["hbm_gbs"]}
    rng = np.random.default_rng(5)
    for name, bs, n in (("bc1_4096", 8, 1 << 20), ("bc7_4096", 16, 1 << 20), ("bc7_16384", 16, 1 << 24)):
        cb = torch.from_numpy(rng.integers(0, 256, (512, bs), dtype=np.uint8)).to(dev)
        idx = torch.from_numpy(rng.integers(0, 512, n, dtype=np.int64).astype(np.int32)).to(dev)
        o = torch.empty(n * bs, dtype=torch.uint8, device=dev)
        best = 1e9
        for _ in range(8):
            enc.timer_start()
            enc.gather_codebook_raw(cb.data_ptr(), 512, bs, idx.data_ptr(), n, o.data_ptr())
            best = min(best, enc.timer_stop())
        ok = bool(torch.equal(o.view(n, bs), cb[idx.long()]))
        gb = n * (4 + bs) / 1e9
        rep[name] = {"blocks": n, "block_bytes": bs, "device_us": best * 1e3, "gbs": gb / (best * 1e-3), "frac_of_hbm_peak": gb / (best * 1e-3) / pk["hbm_gbs"],
                     "correct": ok}
        if n == 1 << 20:
            hcb, hidx = cb.cpu().numpy(), idx.cpu().numpy().view(np.uint32)
            t0 = time.perf_counter()
            for _ in range(5):
                enc.gather_codebook(hcb, hidx)
            rep[name]["host_to_host_ms"] = (time.perf_counter() - t0) / 5 * 1e3
    return rep


def compress_report():
    """the whole VQBCnCompressor::Compress call on the 4096^2 RGB image with the stages outside the GPU core timed
    separately -- Compressonator BC1 pre-pass, device VQ core, K centroid encodes (A16), zstd (the reference prints the sum,
    example/example.cpp:293-304).  Runs through libyuptexture.so (the C++ drop-in) with the Compressonator codec plugin."""
    from yuptexture_b200 import hostapi, synth
    if not hostapi.available():
        return {"unavailable": "libyuptexture.so not built"}
    img = synth.ldr_rgb(TEX, TEX, seed=1000)
    threads = os.cpu_count() or 1
    rep = {"codec": hostapi.codec_name(), "threads": threads}
    for _ in range(2):   # the second call has a warm device arena
        rep["stages_ms"] = hostapi.compress_timed(img, bc_format=1, quality=QUALITY, max_iterations=MAX_ITER, use_zstd=True,
                                                  threads=threads, seed=SEED)
    rep["pre_pass_encoder"] = hostapi.last_encoder_name()
    # the same call with the BC1 pre-pass on the GPU (BCnCompressor::SetGpuEncode(kGpuOn)); host buffers, so
    # bcn_encode_ms includes the 48 MB image upload and the 8 MB block download
    hostapi.set_gpu_bcn_encode(1)
    for _ in range(2):
        rep["stages_ms_gpu_pre_pass"] = hostapi.compress_timed(img, bc_format=1, quality=QUALITY, max_iterations=MAX_ITER, use_zstd=True,
                                                               threads=threads, seed=SEED)
    rep["pre_pass_encoder_gpu"] = hostapi.last_encoder_name()
    hostapi.set_gpu_bcn_encode(-1)
    return rep


def bcn_encode_report(enc, torch, dev, pk):
    """the whole-texture BCn pre-pass (src/vq_bcn_compressor.cpp:94 -> src/bcn_compressor.cpp:45) as a CUDA kernel,
    image and blocks resident in HBM; PSNR of the decoded blocks against Compressonator's on a 1024^2 crop."""
    from yuptexture_b200 import synth
    rep = {}
    for name, fmt, ch, bs in (("bc1_rgb_4096", 1, 3, 8), ("bc3_rgba_4096", 3, 4, 16)):
        img = synth.ldr_rgb(TEX, TEX, seed=1000) if ch == 3 else synth.ldr_rgba(TEX, TEX, seed=1000)
        d_img = torch.from_numpy(img).to(dev)
        d_out = torch.empty(((TEX // 4) ** 2, bs), dtype=torch.uint8, device=dev)
        best = min(enc.bcn_encode_image_raw(d_img.data_ptr(), TEX, TEX, ch, fmt, int(ch == 3), 128, d_out.data_ptr()) for _ in range(5))
        gb = (img.nbytes + d_out.numel()) / 1e9
        rep[name] = {"kernel_ms": best, "gbs": gb / (best * 1e-3), "frac_of_hbm_peak": gb / (best * 1e-3) / pk["hbm_gbs"],
                     "note": "ALU-bound (endpoint search), not HBM-bound"}
        try:
            from oracle import orcbind, refbind
            if refbind.available():
                crop = np.ascontiguousarray(img[:1024, :1024])
                src = synth.image_to_blocks(crop).reshape(-1, 16 * ch)
                if ch == 3:
                    src = src.reshape(-1, 16, 3)[:, :, ::-1].reshape(-1, 48)   # Compressonator reads RGB_888 red/blue-swapped
                ours, _ = enc.bcn_encode_image(crop, fmt)
                t0 = time.perf_counter()
                theirs = refbind.bcn_encode(crop, fmt, quality=1.0, threads=os.cpu_count() or 1)
                t_cmp = time.perf_counter() - t0
                rep[name]["psnr_db_gpu_1024_crop"] = float(synth.psnr_u8(orcbind.decode_bcn(ours, fmt, ch), src))
                rep[name]["psnr_db_compressonator_1024_crop"] = float(synth.psnr_u8(orcbind.decode_bcn(theirs, fmt, ch), src))
                rep[name]["compressonator_ms_1024_crop"] = t_cmp * 1e3
        except Exception as e:   # the checker is optional here
            rep[name]["psnr_check"] = "unavailable: %s" % e
    return rep


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--textures", type=int, default=0, help="textures per step (default: one per GPU = weak scaling; 64 = BASELINE config 5 as stated)")
    ap.add_argument("--placement", default="block", choices=["block", "texture"],
                    help="block: lock-step batch, blocks sharded over the GPUs (headline); texture: whole textures per GPU, no collectives")
    ap.add_argument("--no-extras", action="store_true", help="skip the CPU sample, parity sample, gather and Compress breakdown legs")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
