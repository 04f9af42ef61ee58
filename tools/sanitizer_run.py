"""Reduced-size pass over every kernel family, meant to run under compute-sanitizer:

    compute-sanitizer --tool memcheck  --print-limit 20 python tools/sanitizer_run.py
    compute-sanitizer --tool racecheck --print-limit 20 python tools/sanitizer_run.py --race

Sizes are small (256^2 / 512^2 textures) because the tools slow a launch down by one to two orders of magnitude; the code
paths are the ones the full-size runs take (tensor front-end with resident and with streamed centroid operand, sparse
re-rank, full and delta update, fused tail, palette / pixel-plane / HDR seeding, SAD metric, sampling, lock-step batch).
--race skips the persistent seeding kernels' largest case only (racecheck serialises their grid-wide polls); everything
else is identical.  Prints one line per case and "sanitizer_run: done" at the end.
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yuptexture_b200 import binding, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--race", action="store_true")
    ap.add_argument("--size", type=int, default=512)
    a = ap.parse_args()
    s = a.size
    enc = binding.Encoder(0)

    def case(tag, fn):
        out = fn()
        st = out[2] if isinstance(out, tuple) and len(out) > 2 else None
        extra = " passes %d" % st.lloyd_passes if st is not None and hasattr(st, "lloyd_passes") else ""
        print("ok  %-44s%s" % (tag, extra), flush=True)
        return out

    rgb = synth.ldr_rgb(s, s, seed=1)
    rgba = synth.ldr_rgba(s, s, seed=2)
    bc1 = synth.bc1_encode(rgb)
    bc3 = synth.bc3_encode(rgba)
    n1, n3 = len(bc1), len(bc3)

    # K = 512 (resident centroid operand), Oklab metric, all rows
    cfg = binding.make_config(quality=0.8, max_iterations=8, seed=43)
    cent, idx, _ = case("bc1 K=512 oklab", lambda: enc.encode_bcn(bc1, n1, 1, 3, cfg))
    entries = np.random.default_rng(3).integers(0, 256, (len(cent), 8), dtype=np.uint8)   # BC1-sized codebook entries
    case("gather codebook", lambda: enc.gather_codebook(entries, idx))
    case("bcn encode bc1", lambda: enc.bcn_encode_image(rgb, 1))
    case("bcn encode bc3", lambda: enc.bcn_encode_image(rgba, 3))
    case("compress image (encode + vq)", lambda: enc.compress_image(rgb, 1, cfg)[:3])
    case("bc3 rgba K=512 oklab", lambda: enc.encode_bcn(bc3, n3, 3, 4, cfg))
    # small codebook (single chunk), K = 1024 (streamed operand)
    case("bc1 K=64", lambda: enc.encode_bcn(bc1, n1, 1, 3, binding.make_config(codebook=64, max_iterations=6, seed=5)))
    case("bc1 K=1024 streamed operand", lambda: enc.encode_bcn(bc1, n1, 1, 3, binding.make_config(quality=1.0, max_iterations=4, seed=6)))
    # SAD metric and fast-mode sampling
    case("bc1 K=256 sad", lambda: enc.encode_bcn(bc1, n1, 1, 3, binding.make_config(codebook=256, max_iterations=6, seed=7,
                                                                                      metric=binding.METRIC_SAD)))
    case("bc1 K=256 sampled 0.25", lambda: enc.encode_bcn(bc1, n1, 1, 3, binding.make_config(codebook=256, max_iterations=6, seed=8,
                                                                                               sample_ratio=0.25)))
    # exact (non tensor) assignment path
    case("bc1 K=128 exact assign", lambda: enc.encode_bcn(bc1, n1, 1, 3, binding.make_config(codebook=128, max_iterations=4, seed=9,
                                                                                               assign_mode=binding.ASSIGN_EXACT)))
    # HDR path
    hs = s // 2 if a.race else s
    img = synth.hdr_rgb(hs, hs // 2, seed=11)
    case("hdr K=256", lambda: enc.encode_hdr(img, hs, hs // 2, 3, binding.make_config(codebook=256, max_iterations=6, seed=44)))
    # lock-step batch on one rank (world of one): the batch kernels without a communicator
    if hasattr(enc, "encode_bcn_batch_sharded"):
        tex = [synth.bc1_encode(synth.ldr_rgb(256, 256, seed=20 + t)) for t in range(3)]
        n = len(tex[0])
        try:
            case("batch of 3, 256^2 bc1", lambda: enc.encode_bcn_batch_sharded(tex, n, 0, n, 1, 3,
                                                                                binding.make_config(codebook=128, max_iterations=6, seed=3)))
        except Exception as e:  # a context without a communicator may refuse the sharded entry point
            print("skip batch: %s" % e, flush=True)
    enc.close()
    print("sanitizer_run: done", flush=True)


if __name__ == "__main__":
    main()
