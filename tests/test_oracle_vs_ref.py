"""CPU oracle vs the live, unmodified reference (oracle/_ref/libyupref.so).  Skipped where the
reference library was not built (it needs /root/reference at build time; the built .so travels)."""
import numpy as np
import pytest

from conftest import bits_equal
from yuptexture_b200 import synth


def test_oklab_exhaustive(oracle, reflib):
    """All 2^24 sRGB8 colours: the restatement's FMA pattern and libm calls match the reference TU."""
    c = np.arange(1 << 24, dtype=np.uint32)
    px = np.stack([(c >> 16) & 255, (c >> 8) & 255, c & 255], axis=1).astype(np.uint8).reshape(-1, 48)
    assert bits_equal(reflib.pixels_to_oklab(px, 3), oracle.pixels_to_oklab(px, 3))


@pytest.mark.parametrize("fmt,ch", [(1, 3), (1, 4), (3, 4), (7, 4), (7, 3), (4, 1), (5, 2), (3, 1), (7, 2)])
def test_decode_random(oracle, reflib, fmt, ch):
    rng = np.random.default_rng(100 + fmt * 10 + ch)
    bc = rng.integers(0, 256, (4096, reflib.BLOCK_BYTES[fmt]), dtype=np.uint8)
    if fmt == 7:
        bc[bc[:, 0] == 0, 0] = 1  # reserved encoding: reference output is stale memory
    assert bits_equal(reflib.bcn_decode_blocks(bc.reshape(-1), 256, 256, ch, fmt), oracle.decode_bcn(bc, fmt, ch))


def test_per_block_equals_whole_texture_decode(reflib):
    rng = np.random.default_rng(5)
    bc = rng.integers(0, 256, (64, 8), dtype=np.uint8)
    assert bits_equal(reflib.bcn_decode_each(bc.reshape(-1), 3, 1), reflib.bcn_decode_blocks(bc.reshape(-1), 32, 32, 3, 1))


def test_inverse_and_float_conversions(oracle, reflib):
    rng = np.random.default_rng(9)
    ok = np.empty((4096, 16, 4), np.float32)
    ok[..., 0] = rng.uniform(-0.2, 1.3, ok.shape[:2])
    ok[..., 1:3] = rng.uniform(-0.5, 0.5, ok.shape[:2] + (2,))
    ok[..., 3] = rng.uniform(-0.5, 1.5, ok.shape[:2])
    for ch in (1, 2, 3, 4):
        assert bits_equal(reflib.oklab_to_pixels(ok, ch), oracle.oklab_to_pixels(ok, ch))
        assert bits_equal(reflib.oklab_to_floats(ok, ch), oracle.oklab_to_floats(ok, ch))
        fl = rng.uniform(-5, 300, (2048, 16 * ch)).astype(np.float32)
        assert bits_equal(reflib.floats_to_oklab(fl, ch), oracle.floats_to_oklab(fl, ch))


def test_pipeline_k512(oracle, reflib):
    """quality 0.8 -> K = 512 (the BASELINE configuration), 256x256 BC1, numThreads = 1."""
    img = synth.ldr_rgb(256, 256, seed=3)
    bc = reflib.bcn_encode(img, 1, threads=1)
    px = reflib.bcn_decode_blocks(bc, 256, 256, 3, 1)
    assert bits_equal(px, oracle.decode_bcn(bc, 1, 3))
    c_ref, _, _ = reflib.build_codebook(px, 3, 1, 0.8, 32, 42, threads=1)
    c_orc, _ = oracle.build_codebook(px, 3, 512, 32, 42, threads=1)
    assert bits_equal(c_ref, c_orc)
    i_ref, _ = reflib.quantize_blocks(px, 3, c_ref, threads=1)
    assert bits_equal(i_ref, oracle.quantize_blocks(px, 3, c_ref))
