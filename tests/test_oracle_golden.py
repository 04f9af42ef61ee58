"""CPU oracle (oracle/vq_oracle.cpp) pinned bit-for-bit against the golden vectors generated from
the unmodified reference (tests/golden/make_golden.py).  Runs anywhere g++ exists; no GPU."""
import numpy as np
import pytest

from conftest import bits_equal

FMT = {"bc1": 1, "bc3": 3, "bc4": 4, "bc5": 5, "bc7": 7}


@pytest.mark.parametrize("name,ch", [("bc1", 3), ("bc1", 4), ("bc3", 4), ("bc3", 3), ("bc4", 1), ("bc5", 2),
                                     ("bc5", 4), ("bc7", 4), ("bc7", 3)])
def test_decode_random_blocks(oracle, golden_decode, name, ch):
    got = oracle.decode_bcn(golden_decode[f"{name}_blocks"], FMT[name], ch)
    assert bits_equal(got, golden_decode[f"{name}_px{ch}"])


@pytest.mark.parametrize("name,ch", [("bc1", 3), ("bc3", 4), ("bc7", 4)])
def test_decode_encoder_output(oracle, golden_decode, name, ch):
    got = oracle.decode_bcn(golden_decode[f"{name}_enc_blocks"], FMT[name], ch)
    assert bits_equal(got, golden_decode[f"{name}_enc_px{ch}"])


def test_decode_known_answers(oracle):
    """SURVEY.md Appendix A.3 (measured on the reference's Compressonator 4.3)."""
    def bc1(c0, c1):
        return np.array([c0 & 255, c0 >> 8, c1 & 255, c1 >> 8, 0xE4, 0xE4, 0xE4, 0xE4], np.uint8)
    px = oracle.decode_bcn(bc1(0xF800, 0x0000), 1, 3).reshape(16, 3)
    assert px[:4].tolist() == [[255, 0, 0], [0, 0, 0], [170, 0, 0], [85, 0, 0]]
    px = oracle.decode_bcn(bc1(0xA513, 0x2945), 1, 3).reshape(16, 3)
    assert px[:4].tolist() == [[165, 162, 156], [41, 40, 41], [123, 121, 117], [82, 80, 79]]
    px = oracle.decode_bcn(bc1(0x1234, 0x0123), 1, 4).reshape(16, 4)
    assert px[:4].tolist() == [[16, 69, 165, 255], [0, 36, 24, 255], [10, 58, 118, 255], [5, 47, 71, 255]]
    px = oracle.decode_bcn(bc1(0x0000, 0xF800), 1, 4).reshape(16, 4)
    assert px[:4].tolist() == [[0, 0, 0, 255], [255, 0, 0, 255], [127, 0, 0, 255], [0, 0, 0, 0]]
    px = oracle.decode_bcn(bc1(0x8410, 0x8410), 1, 4).reshape(16, 4)
    assert px[:4].tolist() == [[132, 130, 132, 255]] * 3 + [[0, 0, 0, 0]]
    # BC3 alpha ramps: indices 0..7 for pixels 0..7
    bits = sum(i << (3 * i) for i in range(8))
    def bc3(a0, a1):
        return np.array([a0, a1] + list(bits.to_bytes(6, "little")) + list(bc1(0x2945, 0xA513)), np.uint8)
    px = oracle.decode_bcn(bc3(255, 0), 3, 4).reshape(16, 4)
    assert px[:8, 3].tolist() == [255, 0, 219, 182, 146, 109, 73, 36]
    assert px[2, :3].tolist() == [82, 80, 79] and px[3, :3].tolist() == [123, 121, 117]  # always 4-colour
    px = oracle.decode_bcn(bc3(13, 200), 3, 4).reshape(16, 4)
    assert px[:8, 3].tolist() == [13, 200, 50, 88, 125, 163, 0, 255]


@pytest.mark.parametrize("ch", [1, 2, 3, 4])
def test_colour_conversions(oracle, golden_colour, ch):
    g = golden_colour
    assert bits_equal(oracle.pixels_to_oklab(g[f"px{ch}"], ch), g[f"oklab_from_px{ch}"])
    assert bits_equal(oracle.floats_to_oklab(g[f"fl{ch}"], ch), g[f"oklab_from_fl{ch}"])
    assert bits_equal(oracle.oklab_to_pixels(g["oklab"], ch), g[f"px{ch}_from_oklab"])
    assert bits_equal(oracle.oklab_to_floats(g["oklab"], ch), g[f"fl{ch}_from_oklab"])


def test_oklab_known_answers(oracle):
    px = np.zeros((1, 48), np.uint8)
    px[0, :3] = 255
    px[0, 6:9] = 128
    o = oracle.pixels_to_oklab(px, 3).reshape(16, 4)
    assert o[0, 0] == np.float32(1.0) and abs(o[0, 1]) < 1e-15 and o[0, 3] == 1.0
    assert o[1].tolist() == [0.0, 0.0, 0.0, 1.0]
    assert o[2, 0] == np.float32(0.599870801)


def test_distances(oracle, golden_colour):
    g = golden_colour
    assert bits_equal(oracle.oklab_distance(g["dist_a"], g["dist_b"]), g["dist"])
    assert bits_equal(oracle.block_sad(g["px3"][:64], g["px3"][64:128], 3), g["sad3"])
    assert bits_equal(oracle.block_sad(g["px4"][:64], g["px4"][64:128], 4), g["sad4"])


def test_codebook_size(oracle, golden_colour):
    got = [oracle.codebook_size(float(q)) for q in golden_colour["codebook_size_q"]]
    assert got == golden_colour["codebook_size"].tolist()
    assert oracle.codebook_size(0.8) == 512 and oracle.codebook_size(1.0) == 1024 and oracle.codebook_size(0.5) == 128


LDR_CASES = ["rgb_bc1", "rgba_bc3", "rgba_bc7", "rgba_bc1", "checker_bc1", "const_bc1"]


@pytest.mark.parametrize("name", LDR_CASES)
def test_ldr_pipeline(oracle, golden_pipeline, name):
    g = golden_pipeline
    w, h, fmt, ch, k, seed, iters = (int(v) for v in g[f"{name}_meta"])
    px = oracle.decode_bcn(g[f"{name}_bc"], fmt, ch)
    assert bits_equal(px, g[f"{name}_px"])
    seeds, it0 = oracle.build_codebook(px, ch, k, 0, seed)
    assert it0 == 0 and bits_equal(seeds, g[f"{name}_seeds"])
    cent, _ = oracle.build_codebook(px, ch, k, iters, seed, threads=1)
    assert bits_equal(cent, g[f"{name}_centroids"])
    assert bits_equal(oracle.quantize_blocks(px, ch, cent), g[f"{name}_indices"])


def test_sad_metric_and_sampling(oracle, golden_pipeline):
    g = golden_pipeline
    px = g["sad_px"]
    c, _ = oracle.build_codebook(px, 3, 32, 32, 50, metric=oracle.METRIC_SAD)
    assert bits_equal(c, g["sad_centroids"])
    assert bits_equal(oracle.quantize_blocks(px, 3, c, metric=oracle.METRIC_SAD), g["sad_indices"])
    c, _ = oracle.build_codebook(px, 3, 32, 32, 51, ratio=0.4)
    assert bits_equal(c, g["ratio_centroids"])
    assert bits_equal(oracle.quantize_blocks(px, 3, c), g["ratio_indices"])


@pytest.mark.parametrize("name", ["hdr3", "hdr4"])
def test_hdr_pipeline(oracle, golden_pipeline, name):
    g = golden_pipeline
    w, h, fmt, ch, k, seed, iters = (int(v) for v in g[f"{name}_meta"])
    px = oracle.gather_tiles_f32(g[f"{name}_img"])
    seeds, _ = oracle.build_codebook_hdr(px, ch, k, 0, seed)
    assert bits_equal(seeds, g[f"{name}_seeds"])
    cent, _ = oracle.build_codebook_hdr(px, ch, k, iters, seed, threads=1)
    assert bits_equal(cent, g[f"{name}_centroids"])
    assert bits_equal(oracle.quantize_blocks_hdr(px, ch, cent), g[f"{name}_indices"])


def test_round2_goldens(oracle, golden_pipeline2):
    """pipeline2.npz: the fast-mode sampling prologue at maxIterations = 0 (shuffle + K-Means++ seeds) and a larger HDR case."""
    g = golden_pipeline2
    px = g["ratio_px"]
    for seed in (51, 52, 53):
        ratio, s, k = g[f"ratio_meta_{seed}"]
        c, _ = oracle.build_codebook(px, 3, int(k), 0, int(s), ratio=float(ratio))
        assert bits_equal(c, g[f"ratio_seeds_{seed}"])
    w, h, fmt, ch, k, seed, iters = (int(v) for v in g["hdrL_meta"])
    hp = oracle.gather_tiles_f32(g["hdrL_img"])
    seeds, _ = oracle.build_codebook_hdr(hp, ch, k, 0, seed)
    assert bits_equal(seeds, g["hdrL_seeds"])
    cent, _ = oracle.build_codebook_hdr(hp, ch, k, iters, seed, threads=1)
    assert bits_equal(cent, g["hdrL_centroids"])
    assert bits_equal(oracle.quantize_blocks_hdr(hp, ch, cent), g["hdrL_indices"])
