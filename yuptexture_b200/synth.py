"""Synthetic textures for the BASELINE.json configs.  numpy only, fixed seeds.

C1/C2/C5: per-channel low-frequency sinusoid mix (amplitude 0.4 about 0.5) + uniform noise +-0.075.
C3: same RGB + radial alpha ramp with 10 % of pixels forced to 0/255.
C4: mirrors the recipe of the reference's scripts/hdr_generator.py:9-35 (16x16-px tiles
    uniform[0.1,0.7] per channel, W*H/500 speckles of intensity U[20,100]*(1,1,0.9)).
Adversarial parity textures: constant colour and a two-colour checker.
"""
import numpy as np


def ldr_rgb(width, height, seed=1234):
    rng = np.random.default_rng(seed)
    x = np.arange(width, dtype=np.float32)
    y = np.arange(height, dtype=np.float32)
    img = np.empty((height, width, 3), np.float32)
    for c in range(3):
        acc = np.zeros((height, width), np.float32)
        for _ in range(4):
            fx, fy = rng.uniform(0.5, 6.0, 2) * 2 * np.pi / np.array([width, height])
            ph = rng.uniform(0, 2 * np.pi)
            # sin(fx*x + ph + fy*y) as two outer products
            a, b = (fx * x + ph).astype(np.float32), (fy * y).astype(np.float32)
            acc += np.outer(np.cos(b), np.sin(a)) + np.outer(np.sin(b), np.cos(a))
        img[..., c] = 0.5 + 0.1 * acc
    img += rng.uniform(-0.075, 0.075, img.shape).astype(np.float32)
    return np.clip(np.rint(img * 255.0), 0, 255).astype(np.uint8)


def ldr_rgba(width, height, seed=1234):
    rng = np.random.default_rng(seed + 77)
    rgb = ldr_rgb(width, height, seed)
    y, x = np.mgrid[0:height, 0:width].astype(np.float32)
    r = np.hypot(x - width / 2, y - height / 2) / (0.5 * np.hypot(width, height))
    a = np.clip(np.rint((1.0 - r) * 255.0), 0, 255).astype(np.uint8)
    force = rng.uniform(size=a.shape)
    a[force < 0.05] = 0
    a[force > 0.95] = 255
    return np.concatenate([rgb, a[..., None]], axis=2)


def hdr_rgb(width, height, seed=1234):
    rng = np.random.default_rng(seed)
    bs = 16
    sw, sh = (width + bs - 1) // bs, (height + bs - 1) // bs
    small = rng.uniform(0.1, 0.7, (sh, sw, 3))
    img = np.kron(small, np.ones((bs, bs, 1)))[:height, :width, :].copy()
    n = max(1, (width * height) // 500)
    ys = rng.integers(0, height, n)
    xs = rng.integers(0, width, n)
    inten = rng.uniform(20.0, 100.0, (n, 1))
    img[ys, xs] = inten * np.array([[1.0, 1.0, 0.9]])
    return img.astype(np.float32)


def constant_rgb(width, height, colour=(90, 140, 200)):
    return np.tile(np.array(colour, np.uint8), (height, width, 1))


def checker_rgb(width, height, a=(30, 60, 200), b=(220, 200, 40), cell=8):
    y, x = np.mgrid[0:height, 0:width]
    m = (((x // cell) + (y // cell)) & 1).astype(bool)
    img = np.empty((height, width, 3), np.uint8)
    img[~m] = a
    img[m] = b
    return img


def palette_rgb(width, height, m, seed, jitter=0.0):
    """Texture assembled from m random 4x4 blocks with Zipf-like frequencies (heavy duplicates, m distinct blocks);
    `jitter` of the blocks get +-2 levels of noise.  With m a little above K and few blocks per centroid, Lloyd passes
    AFTER the first one lose clusters, so the empty-cluster reseed runs with non-zero errors (found with
    tools/find_reseed_case.py: 512x256, m = 1100, seed 1, K = 1024 reseeds on every pass)."""
    rng = np.random.default_rng(seed)
    pal = rng.integers(0, 256, (m, 4, 4, 3), dtype=np.uint8)
    nb = (width // 4) * (height // 4)
    p = 1.0 / np.arange(1, m + 1)
    pick = rng.choice(m, nb, p=p / p.sum())
    blocks = pal[pick].astype(np.int16)
    noisy = rng.random(nb) < jitter
    blocks[noisy] += rng.integers(-2, 3, (int(noisy.sum()), 4, 4, 3), dtype=np.int16)
    blocks = np.clip(blocks, 0, 255).astype(np.uint8)
    img = blocks.reshape(height // 4, width // 4, 4, 4, 3).transpose(0, 2, 1, 3, 4).reshape(height, width, 3)
    return np.ascontiguousarray(img)


def image_to_blocks(img):
    """(H, W, ch) -> (n_blocks, 16*ch) 4x4 tiles, edge-clamped, row-major blocks and pixels
    (the gather of reference src/vq_bcn_compressor.cpp:235-252)."""
    h, w, ch = img.shape
    bx, by = (w + 3) // 4, (h + 3) // 4
    ys = np.minimum(np.arange(by * 4), h - 1)
    xs = np.minimum(np.arange(bx * 4), w - 1)
    t = img[ys][:, xs]
    t = t.reshape(by, 4, bx, 4, ch).transpose(0, 2, 1, 3, 4)
    return np.ascontiguousarray(t.reshape(by * bx, 16 * ch))


# ---------------------------------------------------------------------------------------------------
# Minimal BCn encoders (numpy, vectorised) so synthetic BCn inputs exist without the reference's
# Compressonator pre-pass (which is outside the hot path and absent on the GPU box).  Any valid BCn
# stream is a valid input of the VQ path; these use bounding-box endpoints + nearest palette entry.
# ---------------------------------------------------------------------------------------------------

def _pack565(rgb):
    r = (rgb[..., 0].astype(np.uint32) * 31 + 127) // 255
    g = (rgb[..., 1].astype(np.uint32) * 63 + 127) // 255
    b = (rgb[..., 2].astype(np.uint32) * 31 + 127) // 255
    return (r << 11) | (g << 5) | b


def _unpack565(c):
    r, g, b = (c >> 11) & 31, (c >> 5) & 63, c & 31
    return np.stack([(r << 3) | (r >> 2), (g << 2) | (g >> 4), (b << 3) | (b >> 2)], axis=-1).astype(np.int32)


def _bc1_colour_block(rgb, force_four):
    """rgb: (n,16,3) uint8 -> (n,8) uint8 BC1 colour blocks (4-colour mode whenever possible)."""
    n = rgb.shape[0]
    c0 = _pack565(rgb.max(axis=1))
    c1 = _pack565(rgb.min(axis=1))
    swap = c0 < c1
    c0, c1 = np.where(swap, c1, c0), np.where(swap, c0, c1)
    e0, e1 = _unpack565(c0), _unpack565(c1)
    four = (c0 > c1) | force_four
    pal = np.empty((n, 4, 3), np.int32)
    pal[:, 0], pal[:, 1] = e0, e1
    pal[:, 2] = np.where(four[:, None], (2 * e0 + e1) // 3, (e0 + e1) // 2)
    pal[:, 3] = np.where(four[:, None], (e0 + 2 * e1) // 3, 0)
    d = ((rgb[:, :, None, :].astype(np.int32) - pal[:, None, :, :]) ** 2).sum(-1)  # (n,16,4)
    if not force_four:
        d[:, :, 3] = np.where(four[:, None], d[:, :, 3], 1 << 30)  # never pick transparent black
    idx = d.argmin(-1).astype(np.uint32)
    bits = (idx << (2 * np.arange(16, dtype=np.uint32))).sum(axis=1).astype(np.uint32)
    out = np.empty((n, 8), np.uint8)
    out[:, 0], out[:, 1] = c0 & 255, c0 >> 8
    out[:, 2], out[:, 3] = c1 & 255, c1 >> 8
    out[:, 4:8] = bits[:, None] >> (8 * np.arange(4, dtype=np.uint32)) & 255
    return out


def _bc4_block(a):
    """a: (n,16) uint8 -> (n,8) uint8 BC4/BC3-alpha blocks (8-value mode, or flat)."""
    n = a.shape[0]
    a0 = a.max(axis=1).astype(np.int32)
    a1 = a.min(axis=1).astype(np.int32)
    pal = np.empty((n, 8), np.int32)
    pal[:, 0], pal[:, 1] = a0, a1
    eight = a0 > a1
    for i in range(2, 8):
        p8 = ((8 - i) * a0 + (i - 1) * a1 + 3) // 7
        p6 = ((6 - i) * a0 + (i - 1) * a1 + 2) // 5 if i < 6 else (np.zeros(n, np.int32) if i == 6 else np.full(n, 255, np.int32))
        pal[:, i] = np.where(eight, p8, p6)
    idx = np.abs(a[:, :, None].astype(np.int32) - pal[:, None, :]).argmin(-1).astype(np.uint64)
    bits = (idx << (3 * np.arange(16, dtype=np.uint64))).sum(axis=1)
    out = np.empty((n, 8), np.uint8)
    out[:, 0], out[:, 1] = a0, a1
    out[:, 2:8] = (bits[:, None] >> (8 * np.arange(6, dtype=np.uint64))) & 255
    return out


def _chunked(fn, blocks, *args, chunk=1 << 16):
    if blocks.shape[0] <= chunk:
        return fn(blocks, *args)
    return np.concatenate([fn(blocks[i:i + chunk], *args) for i in range(0, blocks.shape[0], chunk)], axis=0)


def bc1_encode(img):
    """(H,W,3) uint8 -> (n_blocks, 8) BC1 blocks."""
    blocks = image_to_blocks(np.ascontiguousarray(img[..., :3])).reshape(-1, 16, 3)
    return _chunked(_bc1_colour_block, blocks, False)


def bc3_encode(img):
    """(H,W,4) uint8 -> (n_blocks, 16) BC3 blocks (alpha block first, then the colour block)."""
    blocks = image_to_blocks(np.ascontiguousarray(img)).reshape(-1, 16, 4)
    return np.concatenate([_bc4_block(blocks[:, :, 3]), _bc1_colour_block(blocks[:, :, :3], force_four=True)], axis=1)


def bc7_encode_mode6(img):
    """(H,W,4) uint8 -> (n_blocks, 16) BC7 mode-6 blocks (one subset, 7777.1 endpoints, 4-bit indices)."""
    blocks = image_to_blocks(np.ascontiguousarray(img)).reshape(-1, 16, 4).astype(np.int32)
    n = blocks.shape[0]
    hi, lo = blocks.max(axis=1), blocks.min(axis=1)
    e = np.stack([lo, hi], axis=1)                        # (n,2,4) 8-bit targets
    p = e[:, :, 0] & 1                                    # one p-bit per endpoint, shared by all channels
    q = np.clip((e - p[:, :, None] + 1) >> 1, 0, 127)     # 7-bit values so that (q<<1|p) ~ e
    full = (q << 1) | p[:, :, None]
    w = np.array([0, 4, 9, 13, 17, 21, 26, 30, 34, 38, 43, 47, 51, 55, 60, 64], np.int32)
    pal = (full[:, 0, None, :] * (64 - w)[None, :, None] + full[:, 1, None, :] * w[None, :, None] + 32) >> 6  # (n,16,4)
    d = ((blocks[:, :, None, :] - pal[:, None, :, :]) ** 2).sum(-1)
    idx = d.argmin(-1)
    # texel 0 is the anchor (its index MSB is implicit 0): swap the endpoints when it lands in the high half
    flip = idx[:, 0] >= 8
    idx = np.where(flip[:, None], 15 - idx, idx).astype(np.uint64)
    q = np.where(flip[:, None, None], q[:, ::-1, :], q)
    p = np.where(flip[:, None], p[:, ::-1], p)
    lo64 = np.full(n, 1 << 6, np.uint64)                  # mode 6: bit 6 set
    hi64 = np.zeros(n, np.uint64)
    pos = 7

    def put(val, nbits):
        nonlocal lo64, hi64, pos
        val = val.astype(np.uint64)
        if pos < 64:
            lo64 |= val << np.uint64(pos)
            if pos + nbits > 64:
                hi64 |= val >> np.uint64(64 - pos)
        else:
            hi64 |= val << np.uint64(pos - 64)
        pos += nbits

    for c in range(4):
        put(q[:, 0, c], 7)
        put(q[:, 1, c], 7)
    put(p[:, 0], 1)
    put(p[:, 1], 1)
    put(idx[:, 0], 3)
    for k in range(1, 16):
        put(idx[:, k], 4)
    assert pos == 128
    out = np.empty((n, 16), np.uint8)
    out[:, :8] = (lo64[:, None] >> (8 * np.arange(8, dtype=np.uint64))) & np.uint64(255)
    out[:, 8:] = (hi64[:, None] >> (8 * np.arange(8, dtype=np.uint64))) & np.uint64(255)
    return out


def psnr_u8(a, b):
    mse = np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)
    return float("inf") if mse == 0 else float(10.0 * np.log10(255.0 ** 2 / mse))
