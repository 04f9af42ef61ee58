"""Generates the committed golden vectors under tests/golden/ from the UNMODIFIED reference
(oracle/_ref/libyupref.so, built by oracle/build_ref.sh from /root/reference).  Run it in the build
container only (the GPU box has no /root/reference); the .npz outputs are committed and are what
the CPU oracle and the CUDA path are pinned against everywhere else.

    python tests/golden/make_golden.py

The reference has no tests or golden vectors of its own, so these are "outputs of
the reference itself run here" in the sense of the parity contract.  Reference calls are made at
numThreads = 1 (its only deterministic accumulation order) with the RNG pinned by the harness.
"""
import os
import sys

import numpy as np

ROOT = os.path.normpath(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
sys.path.insert(0, ROOT)
from oracle import refbind as R  # noqa: E402
from yuptexture_b200 import synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def kat_bc1():
    """Known-answer BC1 blocks of SURVEY.md Appendix A.3 (c0, c1 little-endian, indices E4 E4 E4 E4)."""
    rows = []
    for c0, c1 in [(0xF800, 0x0000), (0xA513, 0x2945), (0x1234, 0x0123), (0x0000, 0xF800), (0x8410, 0x8410)]:
        rows.append([c0 & 255, c0 >> 8, c1 & 255, c1 >> 8, 0xE4, 0xE4, 0xE4, 0xE4])
    return np.array(rows, np.uint8)


def decode_vectors():
    rng = np.random.default_rng(20240601)
    out = {}
    for name, fmt, chans in [("bc1", 1, (3, 4)), ("bc3", 3, (4, 3)), ("bc4", 4, (1,)), ("bc5", 5, (2, 4)),
                             ("bc7", 7, (4, 3))]:
        nb = 1024
        bc = rng.integers(0, 256, (nb, R.BLOCK_BYTES[fmt]), dtype=np.uint8)
        if fmt == 1:
            bc[:5] = kat_bc1()
            bc[5:64, 2:4] = bc[5:64, 0:2]  # c0 == c1 blocks
        if fmt == 3:
            bc[:32, 0] = 255
            bc[:32, 1] = 0
            bc[32:64, 0] = 13
            bc[32:64, 1] = 200
        if fmt == 7:
            # all eight modes well represented; the reserved encoding (first byte 0) is excluded: the
            # reference leaves stale data there and no encoder emits it (decode section).
            for m in range(8):
                sel = slice(m * 96, (m + 1) * 96)
                bc[sel, 0] = (bc[sel, 0] & ~np.uint8((1 << (m + 1)) - 1)) | np.uint8(1 << m)
            bc[bc[:, 0] == 0, 0] = 0x40
        out[f"{name}_blocks"] = bc
        side = 4 * int(np.sqrt(nb))
        for ch in chans:
            out[f"{name}_px{ch}"] = R.bcn_decode_blocks(bc.reshape(-1), side, side, ch, fmt)
    # real encoder output as well (mode statistics of a real texture)
    img = synth.ldr_rgba(64, 64, seed=5)
    for name, fmt, ch in [("bc1", 1, 3), ("bc3", 3, 4), ("bc7", 7, 4)]:
        src = np.ascontiguousarray(img[..., :ch])
        bc = R.bcn_encode(src, fmt, quality=0.2, threads=1)
        out[f"{name}_enc_blocks"] = bc.reshape(-1, R.BLOCK_BYTES[fmt])
        out[f"{name}_enc_px{ch}"] = R.bcn_decode_blocks(bc, 64, 64, ch, fmt)
    np.savez_compressed(os.path.join(OUT, "bc_decode.npz"), **out)


def colour_vectors():
    rng = np.random.default_rng(7)
    out = {}
    for ch in (1, 2, 3, 4):
        px = rng.integers(0, 256, (256, 16 * ch), dtype=np.uint8)
        px[0] = 0
        px[1] = 255
        px[2] = 128
        out[f"px{ch}"] = px
        out[f"oklab_from_px{ch}"] = R.pixels_to_oklab(px, ch)
        fl = rng.uniform(-2.0, 120.0, (128, 16 * ch)).astype(np.float32)
        fl[0] = 0.0
        out[f"fl{ch}"] = fl
        out[f"oklab_from_fl{ch}"] = R.floats_to_oklab(fl, ch)
    ok = np.empty((256, 16, 4), np.float32)
    ok[..., 0] = rng.uniform(-0.1, 1.1, ok.shape[:2])
    ok[..., 1] = rng.uniform(-0.35, 0.35, ok.shape[:2])
    ok[..., 2] = rng.uniform(-0.35, 0.35, ok.shape[:2])
    ok[..., 3] = rng.uniform(-0.2, 1.2, ok.shape[:2])
    ok = ok.reshape(256, 64)
    out["oklab"] = ok
    for ch in (1, 2, 3, 4):
        out[f"px{ch}_from_oklab"] = R.oklab_to_pixels(ok, ch)
        out[f"fl{ch}_from_oklab"] = R.oklab_to_floats(ok, ch)
    a = out["oklab_from_px3"][:96]
    b = out["oklab_from_px4"][:64]
    out["dist_a"], out["dist_b"] = a, b
    out["dist"] = R.oklab_distance(a, b)
    out["sad3"] = R.block_sad(out["px3"][:64], out["px3"][64:128], 3)
    out["sad4"] = R.block_sad(out["px4"][:64], out["px4"][64:128], 4)
    out["codebook_size_q"] = np.array([0.0, 0.1, 0.25, 0.33, 0.5, 0.58, 0.8, 0.92, 1.0, 1.7, -3.0], np.float32)
    out["codebook_size"] = np.array([R.codebook_size(float(q)) for q in out["codebook_size_q"]], np.uint32)
    np.savez_compressed(os.path.join(OUT, "colour.npz"), **out)


def pipeline_vectors():
    out = {}
    cases = [
        ("rgb_bc1", synth.ldr_rgb(128, 128, seed=11), 1, 3, 0.33, 42),
        ("rgba_bc3", synth.ldr_rgba(96, 64, seed=12), 3, 4, 0.25, 43),
        ("rgba_bc7", synth.ldr_rgba(64, 64, seed=13), 7, 4, 0.17, 44),
        ("rgba_bc1", synth.ldr_rgba(64, 64, seed=14), 1, 4, 0.17, 45),
        ("checker_bc1", synth.checker_rgb(64, 64), 1, 3, 0.17, 46),
        ("const_bc1", synth.constant_rgb(32, 32), 1, 3, 0.0, 47),
    ]
    for name, img, fmt, ch, q, seed in cases:
        h, w, _ = img.shape
        bc = R.bcn_encode(np.ascontiguousarray(img), fmt, quality=0.3, threads=1)
        px = R.bcn_decode_blocks(bc, w, h, ch, fmt)
        k = R.codebook_size(q)
        seeds, _, _ = R.build_codebook(px, ch, fmt, q, 0, seed, threads=1)
        cent, _, _ = R.build_codebook(px, ch, fmt, q, 32, seed, threads=1)
        idx, _ = R.quantize_blocks(px, ch, cent, threads=1)
        out[f"{name}_meta"] = np.array([w, h, fmt, ch, k, seed, 32], np.int64)
        out[f"{name}_quality"] = np.array([q], np.float32)
        out[f"{name}_bc"] = bc
        out[f"{name}_px"] = px
        out[f"{name}_seeds"] = seeds      # BuildCodebook(maxIterations=0): seeds after the Oklab round trip
        out[f"{name}_centroids"] = cent
        out[f"{name}_indices"] = idx
    # SAD metric and the fast-mode sampling prologue
    img = synth.ldr_rgb(64, 64, seed=21)
    bc = R.bcn_encode(img, 1, quality=0.3, threads=1)
    px = R.bcn_decode_blocks(bc, 64, 64, 3, 1)
    out["sad_px"] = px
    c, _, _ = R.build_codebook(px, 3, 1, 0.17, 32, 50, metric=R.METRIC_SAD, threads=1)
    out["sad_centroids"] = c
    out["sad_indices"], _ = R.quantize_blocks(px, 3, c, metric=R.METRIC_SAD, threads=1)
    c, _, _ = R.build_codebook(px, 3, 1, 0.17, 32, 51, ratio=0.4, threads=1)
    out["ratio_centroids"] = c
    out["ratio_indices"], _ = R.quantize_blocks(px, 3, c, threads=1)
    # HDR
    for name, ch, seed in [("hdr3", 3, 60), ("hdr4", 4, 61)]:
        img = synth.hdr_rgb(96, 64, seed=seed)
        if ch == 4:
            a = np.random.default_rng(seed).uniform(0, 1, img.shape[:2] + (1,)).astype(np.float32)
            img = np.concatenate([img, a], axis=2)
        px = synth.image_to_blocks(img)
        q = 0.17
        seeds, _, _ = R.build_codebook_hdr(px, ch, q, 0, seed, threads=1)
        cent, _, _ = R.build_codebook_hdr(px, ch, q, 16, seed, threads=1)
        idx, _ = R.quantize_blocks_hdr(px, ch, cent, threads=1)
        out[f"{name}_img"] = img
        out[f"{name}_meta"] = np.array([96, 64, 6, ch, R.codebook_size(q), seed, 16], np.int64)
        out[f"{name}_seeds"] = seeds
        out[f"{name}_centroids"] = cent
        out[f"{name}_indices"] = idx
    np.savez_compressed(os.path.join(OUT, "pipeline.npz"), **out)


def pipeline2_vectors():
    """Round-2 additions (kept in their own file so pipeline.npz stays byte-stable): the fast-mode sampling prologue's
    seed blocks (maxIterations = 0: shuffle + K-Means++ only, bit-exact on the GPU) and a larger HDR case for a
    two-sided 0.05 dB gate."""
    out = {}
    img = synth.ldr_rgb(64, 64, seed=21)
    bc = R.bcn_encode(img, 1, quality=0.3, threads=1)
    px = R.bcn_decode_blocks(bc, 64, 64, 3, 1)
    out["ratio_px"] = px
    for ratio, seed in [(0.4, 51), (0.13, 52), (0.9, 53)]:
        c, _, _ = R.build_codebook(px, 3, 1, 0.17, 0, seed, ratio=ratio, threads=1)
        out[f"ratio_seeds_{seed}"] = c
        out[f"ratio_meta_{seed}"] = np.array([ratio, seed, R.codebook_size(0.17)], np.float64)
    img = synth.hdr_rgb(256, 128, seed=62)
    px = synth.image_to_blocks(img)
    q = 0.33
    out["hdrL_img"] = img
    out["hdrL_meta"] = np.array([256, 128, 6, 3, R.codebook_size(q), 62, 32], np.int64)
    out["hdrL_seeds"], _, _ = R.build_codebook_hdr(px, 3, q, 0, 62, threads=1)
    cent, _, _ = R.build_codebook_hdr(px, 3, q, 32, 62, threads=1)
    out["hdrL_centroids"] = cent
    out["hdrL_indices"], _ = R.quantize_blocks_hdr(px, 3, cent, threads=1)
    np.savez_compressed(os.path.join(OUT, "pipeline2.npz"), **out)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "pipeline2":
        pipeline2_vectors()
        sys.exit(0)
    decode_vectors()
    colour_vectors()
    pipeline_vectors()
    pipeline2_vectors()
    for f in sorted(os.listdir(OUT)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(OUT, f)) // 1024, "KiB")
