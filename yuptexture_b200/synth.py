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
    y, x = np.mgrid[0:height, 0:width].astype(np.float32)
    img = np.empty((height, width, 3), np.float32)
    for c in range(3):
        acc = np.zeros((height, width), np.float32)
        for _ in range(4):
            fx, fy = rng.uniform(0.5, 6.0, 2) * 2 * np.pi / np.array([width, height])
            ph = rng.uniform(0, 2 * np.pi)
            acc += np.sin(fx * x + fy * y + ph).astype(np.float32)
        img[..., c] = 0.5 + 0.4 * acc / 4.0
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
