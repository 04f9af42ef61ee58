"""Drop-in proof against the UNMODIFIED reference (oracle/_ref): tests/cpp/dropin_test.cpp runs
VQBCnCompressor::Compress / CompressHDR through the Compressonator codec plugin (libyupcodec_cmp.so) on BASELINE
config 1 as stated (1024x1024 RGB, BC1, quality 0.8, zstd on, numThreads = host cores) plus BC3 / BC7 / BC7+SAD /
BC6H cases and checks, in C++:
  * the reference's DecompressToBCn decodes OUR .yupt2 stream to the same bytes as ours, and ours decodes a stream
    written by the reference's Compress (src/vq_bcn_compressor.cpp:80-178, :287-361);
  * VQCodebook.entries == the reference's CompressSingleBlock[HDR] on the same pixel centroids, byte for byte (A16).
Here: the reconstructed-texture PSNR of those runs is within 0.05 dB (north_star) of the reference's own
BuildCodebook + QuantizeBlocks on the same decoded blocks and the same seed."""
import os
import subprocess

import numpy as np
import pytest

from yuptexture_b200 import synth

ROOT = os.path.normpath(os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
EXE = os.path.join(ROOT, "build", "dropin_test")
REF = os.path.join(ROOT, "oracle", "_ref", "libyupref.so")
CODEC = os.path.join(ROOT, "yuptexture_b200", "libyupcodec_cmp.so")
PSNR_TOL_DB = 0.05

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dropin(tmp_path_factory):
    if not (os.path.exists(REF) and os.path.exists(CODEC)):
        pytest.skip("needs oracle/_ref/libyupref.so and libyupcodec_cmp.so (built where /root/reference exists)")
    if not os.path.exists(EXE):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "yuptexture_b200", "host")])
    out = tmp_path_factory.mktemp("dropin")
    img = synth.ldr_rgb(1024, 1024, seed=1000)          # BASELINE config 1's texture (the bench generator)
    img.tofile(out / "rgb_1024.u8")
    env = dict(os.environ)
    env.pop("YUPTEXTURE_CODEC", None)
    r = subprocess.run([EXE, str(out), REF, str(out / "rgb_1024.u8")], capture_output=True, text=True, timeout=1200, env=env)
    assert r.returncode == 0 and "DROPIN PASS" in r.stdout, r.stdout[-4000:] + r.stderr[-2000:]
    return out, r.stdout


def test_streams_and_entries_match_reference(dropin):
    _, log = dropin
    assert "Compressonator" in log
    for tag in ("c1_bc1", "c3_bc3", "c3_bc7", "c3_bc7_sad", "hdr_bc6h"):
        line = [l for l in log.splitlines() if l.startswith(tag + ":")]
        assert line, tag
        a, b = line[0].split(":")[1].split("codebook")[0].split("/")
        assert int(a) == int(b) > 0, line[0]


@pytest.mark.parametrize("tag,ch,fmt,quality,metric", [("c1_bc1", 3, 1, 0.8, 1), ("c3_bc3", 4, 3, 0.5, 1), ("c3_bc7", 4, 7, 0.5, 1),
                                                         ("c3_bc7_sad", 4, 7, 0.5, 0)])
def test_compress_psnr_within_tolerance_of_reference(dropin, reflib, tag, ch, fmt, quality, metric):
    out, _ = dropin
    px = np.fromfile(out / f"{tag}_blocks.u8", np.uint8).reshape(-1, 16 * ch)
    cent = np.fromfile(out / f"{tag}_centroids.u8", np.uint8).reshape(-1, 16 * ch)
    idx = np.fromfile(out / f"{tag}_indices.u32", np.uint32)
    threads = os.cpu_count() or 8
    c_ref, _, _ = reflib.build_codebook(px, ch, fmt, quality, 32, 42, metric=metric, threads=threads)
    i_ref, _ = reflib.quantize_blocks(px, ch, c_ref, metric=metric, threads=threads)
    # indices given OUR centroids: the reference's QuantizeBlocks agrees bit for bit
    i_on_ours, _ = reflib.quantize_blocks(px, ch, cent, metric=metric, threads=threads)
    assert np.array_equal(i_on_ours, idx)
    if metric == 0:   # SAD metric: all-integer, bit-exact end to end
        assert np.array_equal(cent, c_ref) and np.array_equal(idx, i_ref)
        return
    p_gpu, p_ref = synth.psnr_u8(cent[idx], px), synth.psnr_u8(c_ref[i_ref], px)
    assert abs(p_gpu - p_ref) <= PSNR_TOL_DB, (p_gpu, p_ref)


def test_compress_hdr_psnr_within_tolerance_of_reference(dropin, reflib):
    out, _ = dropin
    img = np.fromfile(out / "hdr_image.f32", np.float32).reshape(128, 256, 3)
    px = synth.image_to_blocks(img)
    cent = np.fromfile(out / "hdr_centroids.f32", np.float32).reshape(-1, 48)
    idx = np.fromfile(out / "hdr_indices.u32", np.uint32)
    c_ref, _, _ = reflib.build_codebook_hdr(px, 3, 0.5, 16, 42, threads=os.cpu_count() or 8)
    i_ref, _ = reflib.quantize_blocks_hdr(px, 3, c_ref)
    mse_gpu, mse_ref = np.mean((cent[idx] - px) ** 2, dtype=np.float64), np.mean((c_ref[i_ref] - px) ** 2, dtype=np.float64)
    assert abs(10 * np.log10(mse_ref / mse_gpu)) <= PSNR_TOL_DB, (mse_gpu, mse_ref)
