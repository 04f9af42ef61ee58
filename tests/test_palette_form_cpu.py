"""The 20-byte palette form of a BC1 block that the K-Means++ seeding keeps in shared memory
(yuptexture_b200/csrc/convert.cu bc1_palette_planes_kernel, csrc/seed.cu seed_palette_kernel), restated in numpy
and checked on the CPU against the oracle's BC1 decode (reference: BCnCompressor::Decompress as called at
src/vq_bcn_compressor.cpp:142-146): expanding the five words with the PRMT rule must give exactly the 48 pixel
bytes the reference's SAD (src/vq_encoder.cpp:201-208) runs over, channel by channel, so any SAD between two
blocks is the same number in both layouts.  The GPU side of the claim is
tests/test_gpu_pipeline.py::test_bc1_palette_form_seeding_equals_pixel_path."""
import numpy as np


def rgb565(c):
    r5, g6, b5 = (c >> 11) & 31, (c >> 5) & 63, c & 31
    return (r5 << 3) | (r5 >> 2), (g6 << 2) | (g6 >> 4), (b5 << 3) | (b5 >> 2)


def palette_form(bc):
    """n x 8 BC1 bytes -> n x 5 uint32: R, G, B palette words (byte j = entry j), selector nibbles of texels 0-7, 8-15."""
    bc = np.ascontiguousarray(bc, np.uint8).reshape(-1, 8)
    c0 = bc[:, 0].astype(np.int64) | (bc[:, 1].astype(np.int64) << 8)
    c1 = bc[:, 2].astype(np.int64) | (bc[:, 3].astype(np.int64) << 8)
    idx = bc[:, 4:8].copy().view("<u4").reshape(-1).astype(np.int64)
    e0, e1 = rgb565(c0), rgb565(c1)
    four = c0 > c1
    words = []
    for ch in range(3):
        a, b = e0[ch], e1[ch]
        p2 = np.where(four, (2 * a + b) // 3, (a + b) // 2)
        p3 = np.where(four, (a + 2 * b) // 3, 0)  # 3-colour mode: entry 3 is transparent black
        words.append(a | (b << 8) | (p2 << 16) | (p3 << 24))
    for h in range(2):
        x = (idx >> (16 * h)) & 0xFFFF
        x = (x | (x << 8)) & 0x00FF00FF
        x = (x | (x << 4)) & 0x0F0F0F0F
        x = (x | (x << 2)) & 0x33333333
        words.append(x)
    return np.stack(words, axis=1).astype(np.uint32)


def expand(form):
    """PRMT(palette word, 0, selector): byte t of the result = palette byte selected by nibble t.  Returns
    n x 3 x 16 bytes: channel-planar pixels."""
    n = form.shape[0]
    out = np.empty((n, 3, 16), np.uint8)
    sel = np.empty((n, 16), np.int64)
    for t in range(16):
        sel[:, t] = (form[:, 3 + t // 8].astype(np.int64) >> (4 * (t % 8))) & 0xF
    assert sel.max() <= 3
    for ch in range(3):
        pal = form[:, ch].astype(np.int64)
        out[:, ch, :] = (pal[:, None] >> (8 * sel)) & 0xFF
    return out


def test_palette_form_expands_to_the_decoded_pixels(oracle):
    rng = np.random.default_rng(5)
    n = 20000
    bc = rng.integers(0, 256, size=(n, 8), dtype=np.uint8)
    bc[:64, :4] = bc[:64, :1]            # c0 == c1 (3-colour mode, equal endpoints)
    bc[64:128, 4:] = 0xFF                # every texel on entry 3
    px = oracle.decode_bcn(bc, 1, 3).reshape(n, 16, 3)   # reference layout: texel-major RGB
    planar = expand(palette_form(bc))
    assert np.array_equal(planar, px.transpose(0, 2, 1))


def test_sad_is_layout_independent(oracle):
    rng = np.random.default_rng(6)
    bc = rng.integers(0, 256, size=(512, 8), dtype=np.uint8)
    px = oracle.decode_bcn(bc, 1, 3).astype(np.int64)
    planar = expand(palette_form(bc)).reshape(512, 48).astype(np.int64)
    seed = 17
    sad_ref = np.abs(px - px[seed]).sum(axis=1)
    sad_pal = np.abs(planar - planar[seed]).sum(axis=1)
    assert np.array_equal(sad_ref, sad_pal)
