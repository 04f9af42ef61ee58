#!/usr/bin/env python
"""bench.py -- VQ blocks/sec (seeding + Lloyd + assign) on 4096x4096 BC1, the BASELINE.json metric.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one pass of the whole VQ hot path over one synthetic 4096^2 BC1 texture (1,048,576
blocks, K = 512, 32 Lloyd iterations max, Oklab metric): BCn decode -> Oklab features, K-Means++
seeding, Lloyd iterations, centroid conversion, final assignment.  N > 1 (torchrun, one rank per
GPU): weak scaling -- N textures per step in lock step (BASELINE config 5's shape), every texture's blocks
sharded over the N GPUs, seeding texture-parallel (texture t on rank t % N), and ONE NCCL all-reduce per
Lloyd pass carrying the K x 65 fixed-point sums/counts of all N textures (the exchange step north_star names).

`value`  : blocks/s with the BCn bytes already resident in HBM and outputs left in HBM.
`e2e`    : the same through the C ABI with HOST (pinned) buffers: H2D of the BCn bytes and D2H of the
           centroids + indices inside the timed region.
--impl reference : the unmodified reference (oracle/_ref, else the CPU port) on the host cores, on
           a bounded sample (a crop of the same texture) per step.
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
CPU_SAMPLE = 1024   # the reference is timed on a CPU_SAMPLE^2 crop (bounded sample)


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


def cpu_sample_pixels(bc):
    """Decoded pixel blocks of the top-left CPU_SAMPLE^2 crop of the bench texture."""
    from oracle import orcbind
    bx = TEX // 4
    rows = np.arange(CPU_SAMPLE // 4)[:, None] * bx + np.arange(CPU_SAMPLE // 4)[None, :]
    return orcbind.decode_bcn(bc[rows.reshape(-1)], 1, 3)


def run_reference(args, rank, world):
    if rank != 0:
        return
    bc = make_texture(1000)
    px = cpu_sample_pixels(bc)
    threads = os.cpu_count() or 1
    for _ in range(args.warmup):
        cpu_reference_step(px, threads)
    t0 = time.perf_counter()
    kind = "port"
    for _ in range(args.steps):
        _, kind = cpu_reference_step(px, threads)[:2]
    dt = (time.perf_counter() - t0) / max(1, args.steps)
    value = len(px) / dt
    sample = f"{CPU_SAMPLE}x{CPU_SAMPLE} crop ({len(px)} blocks) of the 4096^2 BC1 bench texture per step, K=512, 32 iterations max"
    print(json.dumps({
        "impl": "reference", "metric": "vq_blocks_per_sec", "value": value, "unit": "blocks/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "BC1 VQ compress of a synthetic 4096x4096 texture (config 2): K=512, maxIterations=32, Oklab",
                   "note": "CPU reference timed on a bounded sample"},
        "cpu_baseline": {"value": value, "unit": "blocks/s", "cores": threads if kind == "reference" else 1,
                         "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "blocks/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from yuptexture_b200 import binding

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    enc = binding.Encoder(local_rank)
    k = binding.codebook_size(QUALITY)
    n = (TEX // 4) ** 2
    n_tex = world  # weak scaling: one texture's worth of blocks per GPU
    cfg = binding.make_config(quality=QUALITY, max_iterations=MAX_ITER, seed=SEED)

    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            uid = torch.from_numpy(enc.comm_unique_id().copy())
        uid = uid.to(dev)
        dist.broadcast(uid, 0)
        enc.comm_init(uid.cpu().numpy(), world, rank)

    # synthetic inputs (fixed seeds): texture t is generated on rank t % world and broadcast, every rank holds all of them
    textures = []
    for t in range(n_tex):
        if world == 1:
            textures.append(make_texture(1000 + t))
            continue
        buf = torch.empty(n * 8, dtype=torch.uint8, device=dev)
        if t % world == rank:
            buf.copy_(torch.from_numpy(make_texture(1000 + t).reshape(-1).copy()))
        dist.broadcast(buf, t % world)
        textures.append(buf.cpu().numpy().reshape(n, 8))
    pinned_in = [torch.from_numpy(t.reshape(-1).copy()).pin_memory() for t in textures]
    dev_in = [p.to(dev) for p in pinned_in]
    from yuptexture_b200 import sharding
    first, n_local = sharding.shard_range(n, world, rank)
    dev_cent = [torch.empty(k * 48, dtype=torch.uint8, device=dev) for _ in range(n_tex)]
    dev_idx = [torch.empty(max(1, n_local), dtype=torch.int32, device=dev) for _ in range(n_tex)]
    host_cent = [torch.empty(k * 48, dtype=torch.uint8).pin_memory() for _ in range(n_tex)]
    host_idx = [torch.empty(max(1, n_local), dtype=torch.int32).pin_memory() for _ in range(n_tex)]

    def step(resident):
        if world == 1:
            src = dev_in[0].data_ptr() if resident else pinned_in[0].numpy()
            oc = dev_cent[0].data_ptr() if resident else host_cent[0].numpy()
            oi = dev_idx[0].data_ptr() if resident else host_idx[0].numpy().view(np.uint32)
            _, _, st = enc.encode_bcn(src, n, 1, 3, cfg, oc, oi)
            return [st]
        # N GPUs: N textures in lock step, every texture's blocks sharded over the N GPUs; one all-reduce per pass
        srcs = [d.data_ptr() for d in dev_in] if resident else [p.numpy() for p in pinned_in]
        ocs = [d.data_ptr() for d in dev_cent] if resident else [h.numpy() for h in host_cent]
        ois = [d.data_ptr() for d in dev_idx] if resident else [h.numpy().view(np.uint32) for h in host_idx]
        _, _, st = enc.encode_bcn_batch_sharded(srcs, n, first, n_local, 1, 3, cfg, ocs, ois)
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

    total_blocks = n * n_tex
    value = total_blocks * args.steps / (ms_dev * 1e-3)
    e2e = total_blocks * args.steps / (ms_e2e * 1e-3)

    # one more real step (all ranks: it contains collectives) with the front-end launches bracketed by events
    enc.profile(True)
    step(True)
    ms_fe, n_fe = enc.profile_read()
    enc.profile(False)

    out = None
    if rank == 0:
        pk = peaks()
        # dominant kernel: the tcgen05 assignment front-end (assign_tc_kernel), every launch of it inside one
        # more real step bracketed by CUDA events on the library's stream (yupvq_profile)
        ms_a = ms_fe / max(1, n_fe)
        blocks_per_launch = n_local * n_tex if world > 1 else n   # the lock-step batch launches the front-end once for all textures
        flop_alg = 2.0 * blocks_per_launch * k * 64          # 2 K D per block per pass
        split = 3                                            # xh.ch + xh.cl + xl.ch bf16 partial products
        achieved = split * flop_alg / (ms_a * 1e-3) / 1e12 if ms_a > 0 else 0.0
        traffic = None
        import glob
        tfiles = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_traffic.json")))
        tpath = tfiles[-1] if tfiles else ""
        if world == 1 and tpath:  # DRAM bytes per launch of the same kernel on the same config, from the committed ncu capture
            traffic = json.load(open(tpath))["assign_tc_kernel"]["dram_bytes_per_launch"]
        roofline = {"bound": "tensor", "kernel": "assign_tc_kernel (N x K x 64 contraction as 3 bf16 partial products + argmax epilogue)",
                    "achieved": achieved, "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s",
                    "frac": achieved / pk["bf16_tflops_sustained"], "traffic": traffic,
                    "peak_source": pk["src"] + " bf16 sustained (kernel timed inside a step)",
                    "ms_per_launch": ms_a, "launches_timed": n_fe, "flop_per_launch_algorithmic": flop_alg, "bf16_split": split,
                    "achieved_unsplit": flop_alg / (ms_a * 1e-3) / 1e12 if ms_a > 0 else 0.0,
                    "share_of_step": ms_fe / (ms_dev / args.steps) if ms_dev > 0 else None}
        # bounded CPU sample of the same workload on this box's host cores
        px = cpu_sample_pixels(textures[0])
        threads = os.cpu_count() or 1
        dt, kind, cent_ref, idx_ref = cpu_reference_step(px, threads)
        solo = binding.Encoder(local_rank)   # no communicator: the parity sample is a single-GPU call on rank 0 only
        parity = parity_report(solo, px, cent_ref, idx_ref)
        solo.close()
        cpu = {"value": len(px) / dt, "unit": "blocks/s", "cores": threads if kind == "reference" else 1, "kind": kind,
               "sample": f"{CPU_SAMPLE}x{CPU_SAMPLE} crop ({len(px)} blocks) of the bench texture, one pass, "
                         f"BuildCodebook+QuantizeBlocks {dt:.2f} s"}
        bytes_in = n * 8 * n_tex
        bytes_out = (k * 48 + n_local * 4) * n_tex
        phases = ("h2d", "decode", "seed", "lloyd", "finalize", "quantize", "d2h")
        out = {
            "metric": "vq_blocks_per_sec", "value": value, "unit": "blocks/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"BC1 VQ compress of {n_tex} synthetic 4096x4096 texture(s) (config 2 per GPU): "
                                   "seeding+Lloyd+assign, K=512, maxIterations=32, Oklab metric",
                       "blocks_per_texture": n, "textures": n_tex, "codebook": k,
                       "parallelism": "1 GPU" if world == 1 else f"{n_tex} textures in lock step, blocks sharded over {world} GPUs, seeding texture-parallel, "
                                      f"one NCCL all-reduce of {n_tex} x K x 65 int64 per pass",
                       "l2": "no explicit flush: each step streams a 330 MB working set (features + pixel planes) > 126 MB L2",
                       "lloyd_passes": last.lloyd_passes, "ms_per_texture": ms_dev / args.steps / n_tex,
                       "phases_ms": {p: getattr(last, "ms_" + p) for p in phases},
                       "wall_ms_per_step": wall_dev / args.steps},
            "e2e": {"value": e2e, "unit": "blocks/s", "h2d_bytes_per_step": bytes_in, "d2h_bytes_per_step": bytes_out,
                    "ms_per_step": ms_e2e / args.steps, "wall_ms_per_step": wall_e2e / args.steps},
            "gpu_launches": launches,
            "roofline": roofline,
            "cpu_baseline": cpu,
            "parity": parity,
            "clocks": clocks.summary(),
        }
        print(json.dumps(out))
    enc.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
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
