"""N4: the GPU BCn encoders' block code, compiled as host C++, gated against Compressonator (oracle/_ref) on
decoded PSNR -- the bar VERDICT r01 set for the row.  Reference: src/bcn_compressor.cpp:45-116.
The GPU tests (tests/test_gpu_bcn_encode.py) check that the CUDA kernel writes exactly these bytes."""
import numpy as np
import pytest

import bcnenc_host
from yuptexture_b200 import synth


def _swap(src, ch):
    return src.reshape(-1, 16, ch)[:, :, [2, 1, 0] + ([3] if ch == 4 else [])].reshape(src.shape)


def _cases():
    rgb = synth.ldr_rgb(512, 512, seed=1000)
    rgba = synth.ldr_rgba(512, 512, seed=1000)
    return {
        # name: (image, format, decode channels, compare against the red/blue-swapped source?)
        "bc1_rgb": (rgb, 1, 3, True),            # Compressonator reads RGB_888 swapped (flipRGB == false on the VQ path)
        "bc3_rgba": (rgba, 3, 4, False),
        "bc4_grey": (rgb[:, :, :1].copy(), 4, 1, False),
        "bc4_alpha_outliers": (rgba[:, :, 3:].copy(), 4, 1, False),   # smooth ramp with forced 0 / 255 texels: 6-value mode
    }


@pytest.mark.parametrize("name", ["bc1_rgb", "bc3_rgba", "bc4_grey", "bc4_alpha_outliers"])
def test_decoded_psnr_at_least_compressonator(name, oracle, reflib):
    img, fmt, ch, swapped = _cases()[name]
    src = synth.image_to_blocks(img).reshape(-1, 16 * ch)
    if swapped:
        src = _swap(src, ch)
    ours = oracle.decode_bcn(bcnenc_host.encode(img, fmt), fmt, ch)
    cmp_ = oracle.decode_bcn(reflib.bcn_encode(img, fmt, quality=1.0), fmt, ch)
    p_ours, p_cmp = synth.psnr_u8(ours, src), synth.psnr_u8(cmp_, src)
    assert p_ours >= p_cmp, (name, p_ours, p_cmp)     # measured: +0.1 .. +0.4 dB, lossless where Compressonator is


def test_punch_through_alpha_matches_threshold(oracle, reflib):
    rgba = synth.ldr_rgba(256, 256, seed=5)
    src = synth.image_to_blocks(rgba).reshape(-1, 16, 4)
    opaque = src[:, :, 3] >= 128
    ours = oracle.decode_bcn(bcnenc_host.encode(rgba, 1), 1, 4).reshape(-1, 16, 4)
    cmp_ = oracle.decode_bcn(reflib.bcn_encode(rgba, 1, quality=1.0), 1, 4).reshape(-1, 16, 4)
    assert np.array_equal(ours[:, :, 3] > 0, opaque) and np.array_equal(cmp_[:, :, 3] > 0, opaque)

    def mse(d):
        return ((d[:, :, :3].astype(np.float64) - src[:, :, :3]) ** 2)[opaque].mean()
    assert mse(ours) <= mse(cmp_)


def test_flat_and_two_colour_blocks_exact_where_representable(oracle):
    for img in (synth.constant_rgb(64, 64), synth.checker_rgb(64, 64)):     # 8x8 cells: every block is flat
        src = _swap(synth.image_to_blocks(img).reshape(-1, 48), 3)
        got = oracle.decode_bcn(bcnenc_host.encode(img, 1), 1, 3)
        assert np.abs(got.astype(int) - src.astype(int)).max() <= 1
    # 565-representable two-colour blocks come back exactly
    a, b = (8, 4, 8), (255, 251, 247)
    img = np.empty((8, 8, 3), np.uint8)
    img[:] = a
    img[::2, ::2] = b
    src = synth.image_to_blocks(img).reshape(-1, 48)
    assert np.array_equal(oracle.decode_bcn(bcnenc_host.encode(img, 1, swap_rb=False), 1, 3), src)


def test_ragged_edges_and_channel_expansion(oracle):
    img = synth.ldr_rgb(70, 50, seed=3)
    bc = bcnenc_host.encode(img, 1)
    assert bc.shape == (18 * 13, 8)
    src = _swap(synth.image_to_blocks(img).reshape(-1, 48), 3)
    assert synth.psnr_u8(oracle.decode_bcn(bc, 1, 3), src) > 26.0
    # BC5 of a 2-channel source: first channel -> red block, second -> green block
    rg = synth.ldr_rgb(64, 64, seed=9)[:, :, :2].copy()
    dec = oracle.decode_bcn(bcnenc_host.encode(rg, 5), 5, 4).reshape(-1, 16, 4)
    s2 = synth.image_to_blocks(rg).reshape(-1, 16, 2)
    assert synth.psnr_u8(dec[:, :, :2], s2) > 38.0      # 64-texel sine periods: steep ramps inside a block
