"""N4 on the GPU: the BCn encode kernel (yuptexture_b200/csrc/bcn_encode.cu) through the C ABI.
Reference stage: BCnCompressor::Compress, src/bcn_compressor.cpp:45-116 (called at src/vq_bcn_compressor.cpp:94)."""
import numpy as np
import pytest

import bcnenc_host
from yuptexture_b200 import binding, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def enc():
    e = binding.Encoder(0)
    yield e
    e.close()


def _swap(src, ch):
    return src.reshape(-1, 16, ch)[:, :, [2, 1, 0] + ([3] if ch == 4 else [])].reshape(src.shape)


@pytest.mark.parametrize("fmt,ch,size", [(1, 3, (512, 512)), (1, 4, (256, 256)), (3, 4, (512, 512)), (4, 1, (256, 256)), (5, 2, (256, 256)),
                                         (1, 3, (70, 50)), (3, 4, (33, 47)), (5, 3, (64, 64)), (4, 2, (64, 64))])
def test_kernel_bytes_equal_host_compiled_block_code(enc, fmt, ch, size):
    """Same source, same rounding (no FMA contraction on either side): the kernel's blocks are the host build's blocks."""
    w, h = size
    img = (synth.ldr_rgba(w, h, seed=7) if ch == 4 else synth.ldr_rgb(w, h, seed=7)[:, :, :ch]).copy()
    got, _ = enc.bcn_encode_image(img, fmt)
    want = bcnenc_host.encode(img, fmt)
    assert got.shape == want.shape
    bad = np.nonzero((got != want).any(1))[0]
    assert bad.size == 0, (bad[:8], got[bad[:2]], want[bad[:2]])


def test_flat_checker_and_extreme_blocks(enc, oracle):
    for img in (synth.constant_rgb(64, 64), synth.checker_rgb(64, 64), np.zeros((16, 16, 4), np.uint8), np.full((16, 16, 4), 255, np.uint8)):
        ch = img.shape[2]
        for fmt in ((1, 3) if ch == 4 else (1,)):
            got, _ = enc.bcn_encode_image(img, fmt)
            assert np.array_equal(got, bcnenc_host.encode(img, fmt))
            src = synth.image_to_blocks(img).reshape(-1, 16 * ch)
            if ch == 3:
                src = _swap(src, 3)
            dec = oracle.decode_bcn(got, fmt, ch).astype(int)
            if ch == 4 and fmt == 1 and img[0, 0, 3] == 0:
                assert (dec == 0).all()          # punch-through: transparent black
            else:
                assert np.abs(dec - src.astype(int)).max() <= 1


def test_config2_texture_psnr_vs_compressonator(enc, oracle, reflib):
    """4096^2 is the bench size; the gate runs on a 1024^2 crop so that Compressonator finishes in seconds."""
    img = synth.ldr_rgb(1024, 1024, seed=1000)
    src = _swap(synth.image_to_blocks(img).reshape(-1, 48), 3)
    got, ms = enc.bcn_encode_image(img, 1)
    p_gpu = synth.psnr_u8(oracle.decode_bcn(got, 1, 3), src)
    p_cmp = synth.psnr_u8(oracle.decode_bcn(reflib.bcn_encode(img, 1, quality=1.0), 1, 3), src)
    assert p_gpu >= p_cmp, (p_gpu, p_cmp)
    assert ms < 5.0


def test_compress_image_equals_encode_then_vq(enc):
    """yupvq_compress_image keeps the blocks in HBM; results equal the two separate calls."""
    img = synth.ldr_rgb(512, 512, seed=11)
    cfg = binding.make_config(codebook=128, max_iterations=8, seed=3)
    bc, _ = enc.bcn_encode_image(img, 1)
    cent_a, idx_a, _ = enc.encode_bcn(bc, len(bc), 1, 3, cfg)
    cent_b, idx_b, st, ms, bc_b = enc.compress_image(img, 1, cfg, want_bc=True)
    assert np.array_equal(bc, bc_b) and np.array_equal(cent_a, cent_b) and np.array_equal(idx_a, idx_b)
    assert st.lloyd_passes >= 1 and ms > 0.0


@pytest.mark.parametrize("fmt,ch", [(1, 3), (3, 4)])
def test_misaligned_device_output_is_staged(enc, fmt, ch):
    """a device output pointer that is not aligned to the block size (a slice of a larger buffer) must not reach
    the kernel's uint2 / uint4 stores; the call stages it and the bytes equal the aligned call's."""
    import torch
    img = (synth.ldr_rgba(128, 64, seed=5) if ch == 4 else synth.ldr_rgb(128, 64, seed=5)).copy()
    want, _ = enc.bcn_encode_image(img, fmt)
    d_img = torch.from_numpy(img).cuda()
    d_buf = torch.full((want.size + 64,), 0xAB, dtype=torch.uint8, device="cuda")
    off = 4 if fmt == 1 else 8                      # multiple of 4 (fine for any copy), not of the block size
    assert (d_buf.data_ptr() + off) % want.shape[1] != 0
    torch.cuda.synchronize()
    enc.bcn_encode_image_raw(d_img.data_ptr(), 128, 64, ch, fmt, int(ch == 3), 128, d_buf.data_ptr() + off)
    torch.cuda.synchronize()
    got = d_buf.cpu().numpy()
    assert np.array_equal(got[off:off + want.size].reshape(want.shape), want)
    assert (got[:off] == 0xAB).all() and (got[off + want.size:] == 0xAB).all()


def test_full_size_properties(enc):
    """BASELINE config 2 size: every block decodes close to its source texels; device-resident input and output."""
    import torch
    img = synth.ldr_rgb(4096, 4096, seed=1000)
    d_img = torch.from_numpy(img).cuda()
    d_out = torch.empty((1 << 20, 8), dtype=torch.uint8, device="cuda")
    ms = enc.bcn_encode_image_raw(d_img.data_ptr(), 4096, 4096, 3, 1, 1, 128, d_out.data_ptr())
    torch.cuda.synchronize()
    bc = d_out.cpu().numpy()
    dec = enc.decode_bcn(bc, 1, 3)
    src = _swap(synth.image_to_blocks(img).reshape(-1, 48), 3)
    err = ((dec.astype(np.float32) - src) ** 2).reshape(len(src), -1).mean(1)
    assert synth.psnr_u8(dec, src) > 29.5 and err.max() < 400.0      # no block far off (noise amplitude +-19 levels)
    # endpoints in 4-colour order or equal (this texture has no alpha)
    c0 = bc[:, 0].astype(np.uint32) | (bc[:, 1].astype(np.uint32) << 8)
    c1 = bc[:, 2].astype(np.uint32) | (bc[:, 3].astype(np.uint32) << 8)
    assert (c0 >= c1).all()
    assert ms < 2.0
