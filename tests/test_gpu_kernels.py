"""Kernel-level parity on the B200, through the C ABI: every stage of the path against the golden
vectors generated from the unmodified reference and against the CPU oracle on seeded inputs.
Integer/byte/index results are bit-exact; the only tolerances are on float paths that need pow()
on arbitrary arguments (stated at each assert)."""
import numpy as np
import pytest

from conftest import bits_equal
from yuptexture_b200 import synth

pytestmark = pytest.mark.gpu

FMT = {"bc1": 1, "bc3": 3, "bc4": 4, "bc5": 5, "bc7": 7}


@pytest.fixture(scope="module")
def enc():
    from yuptexture_b200 import binding
    e = binding.Encoder(0)
    yield e
    e.close()


def ulp_diff(a, b):
    a = np.ascontiguousarray(a, np.float32).view(np.int32).astype(np.int64)
    b = np.ascontiguousarray(b, np.float32).view(np.int32).astype(np.int64)
    a = np.where(a < 0, -(a & 0x7fffffff), a)
    b = np.where(b < 0, -(b & 0x7fffffff), b)
    return np.abs(a - b)


# ---- K1: decode -------------------------------------------------------------------------------------

@pytest.mark.parametrize("name,ch", [("bc1", 3), ("bc1", 4), ("bc3", 4), ("bc3", 3), ("bc4", 1), ("bc5", 2),
                                     ("bc5", 4), ("bc7", 4), ("bc7", 3)])
def test_decode_golden(enc, golden_decode, name, ch):
    assert bits_equal(enc.decode_bcn(golden_decode[f"{name}_blocks"], FMT[name], ch), golden_decode[f"{name}_px{ch}"])


@pytest.mark.parametrize("name,ch", [("bc1", 3), ("bc3", 4), ("bc7", 4)])
def test_decode_golden_encoder_output(enc, golden_decode, name, ch):
    assert bits_equal(enc.decode_bcn(golden_decode[f"{name}_enc_blocks"], FMT[name], ch),
                      golden_decode[f"{name}_enc_px{ch}"])


@pytest.mark.parametrize("fmt,ch,n", [(1, 3, 1 << 20), (1, 4, 100003), (3, 4, 1 << 18), (7, 4, 1 << 18), (7, 3, 77777),
                                      (4, 1, 5001), (5, 2, 5001), (1, 1, 999), (7, 2, 999), (3, 2, 1001)])
def test_decode_vs_oracle_random(enc, oracle, fmt, ch, n):
    rng = np.random.default_rng(fmt * 100 + ch)
    bc = rng.integers(0, 256, (n, oracle.BLOCK_BYTES[fmt]), dtype=np.uint8)
    assert bits_equal(enc.decode_bcn(bc, fmt, ch), oracle.decode_bcn(bc, fmt, ch))


def test_decode_empty_and_single(enc, oracle):
    assert enc.decode_bcn(np.zeros((0, 8), np.uint8), 1, 3).shape == (0, 48)
    bc = np.array([[0x00, 0xF8, 0x00, 0x00, 0xE4, 0xE4, 0xE4, 0xE4]], np.uint8)
    assert enc.decode_bcn(bc, 1, 3).reshape(16, 3)[:4].tolist() == [[255, 0, 0], [0, 0, 0], [170, 0, 0], [85, 0, 0]]


# ---- K1: colour conversion ----------------------------------------------------------------------------

@pytest.mark.parametrize("ch", [1, 2, 3, 4])
def test_pixels_to_oklab_golden(enc, golden_colour, ch):
    assert bits_equal(enc.pixels_to_oklab(golden_colour[f"px{ch}"], ch), golden_colour[f"oklab_from_px{ch}"])


def test_pixels_to_oklab_exhaustive(enc, oracle):
    """All 2^24 sRGB8 colours, bit-exact (device cbrt restates glibc's; sRGB table comes from host libm)."""
    c = np.arange(1 << 24, dtype=np.uint32)
    px = np.stack([(c >> 16) & 255, (c >> 8) & 255, c & 255], axis=1).astype(np.uint8).reshape(-1, 48)
    assert bits_equal(enc.pixels_to_oklab(px, 3), oracle.pixels_to_oklab(px, 3))


@pytest.mark.parametrize("ch", [1, 2, 3, 4])
def test_oklab_to_pixels_golden(enc, golden_colour, ch):
    # uint8 outputs: CUDA pow vs glibc pow can only matter within ~1e-13 of a rounding boundary
    assert bits_equal(enc.oklab_to_pixels(golden_colour["oklab"], ch), golden_colour[f"px{ch}_from_oklab"])


@pytest.mark.parametrize("ch", [1, 2, 3, 4])
def test_float_conversions_golden(enc, golden_colour, ch):
    """pow() on arbitrary doubles: CUDA's pow is not glibc's; tolerance = 1 float ulp (or 1e-12 absolute
    for the cancelling near-zero a/b of greys)."""
    got = enc.floats_to_oklab(golden_colour[f"fl{ch}"], ch)
    ref = golden_colour[f"oklab_from_fl{ch}"]
    ok = (ulp_diff(got, ref) <= 1) | (np.abs(got.astype(np.float64) - ref) <= 1e-12)
    assert ok.all(), f"{(~ok).sum()} values off by more than 1 ulp"
    big = np.abs(ref) > 1e-9  # the a/b of greys cancel to ~1e-17: any last-bit pow difference shows there
    assert (ulp_diff(got, ref)[big] == 0).mean() > 0.999
    got = enc.oklab_to_floats(golden_colour["oklab"], ch)
    ref = golden_colour[f"fl{ch}_from_oklab"]
    ok = (ulp_diff(got, ref) <= 1) | (np.abs(got.astype(np.float64) - ref) <= 1e-9)
    assert ok.all()


def test_gather_tiles(enc):
    img = synth.hdr_rgb(70, 37, seed=3)  # ragged edges: clamp path
    assert bits_equal(enc.gather_tiles(img), synth.image_to_blocks(img))


# ---- K3: assignment ------------------------------------------------------------------------------------

def test_distance_golden_via_assign(enc, golden_colour):
    """errors[i] must equal the reference's OklabBlockDistanceSq_SIMD bit for bit."""
    a, b, d = golden_colour["dist_a"], golden_colour["dist_b"], golden_colour["dist"]
    idx, err, _ = enc.assign(a, b, assign_mode=1)
    assert bits_equal(err, d.min(axis=1))
    assert np.array_equal(idx, d.argmin(axis=1).astype(np.uint32))


@pytest.mark.parametrize("n,k", [(100000, 512), (4097, 1024), (33, 16), (1, 7), (5000, 300)])
def test_assign_exact_vs_oracle(enc, oracle, n, k):
    rng = np.random.default_rng(n + k)
    px = rng.integers(0, 256, (n, 48), dtype=np.uint8)
    pc = rng.integers(0, 256, (k, 48), dtype=np.uint8)
    f, c = oracle.pixels_to_oklab(px, 3), oracle.pixels_to_oklab(pc, 3)
    i_ref, e_ref = oracle.assign(f, c)
    i_gpu, e_gpu, _ = enc.assign(f, c, assign_mode=1)
    assert bits_equal(i_gpu, i_ref) and bits_equal(e_gpu, e_ref)


def test_assign_ties_lowest_index(enc, oracle):
    rng = np.random.default_rng(5)
    c = oracle.pixels_to_oklab(rng.integers(0, 256, (8, 48), dtype=np.uint8), 3)
    cent = np.concatenate([c, c, c[:4]])          # duplicates: exact ties
    feat = np.concatenate([c, c[::-1]])           # zero distance to several centroids
    idx, err, _ = enc.assign(feat, cent, assign_mode=1)
    i_ref, e_ref = oracle.assign(feat, cent)
    assert bits_equal(idx, i_ref) and bits_equal(err, e_ref) and idx.max() < 8 and (err == 0).all()


# ---- K3: tensor engine (tcgen05 front-end + exact re-rank) must equal the exact engine bit for bit ---------

@pytest.mark.parametrize("n,k", [(100000, 512), (4097, 1024), (8192 + 77, 300), (1, 7), (5000, 64), (129, 2000)])
def test_assign_tensor_random_pixels(enc, oracle, n, k):
    """Uniform-random pixel blocks: dense near-ties, long candidate lists and the overflow scan."""
    rng = np.random.default_rng(n * 7 + k)
    f = oracle.pixels_to_oklab(rng.integers(0, 256, (n, 48), dtype=np.uint8), 3)
    c = oracle.pixels_to_oklab(rng.integers(0, 256, (k, 48), dtype=np.uint8), 3)
    i_ref, e_ref = oracle.assign(f, c)
    i_gpu, e_gpu, _ = enc.assign(f, c, assign_mode=2)
    assert bits_equal(i_gpu, i_ref) and bits_equal(e_gpu, e_ref)


@pytest.mark.parametrize("k,ch", [(512, 3), (1024, 3), (256, 4), (128, 4)])
def test_assign_tensor_texture(enc, oracle, k, ch):
    """A real texture against a trained codebook: the short-list path (1-2 candidates per block)."""
    img = synth.ldr_rgb(512, 512, seed=k) if ch == 3 else synth.ldr_rgba(512, 512, seed=k)
    px = synth.image_to_blocks(img)
    cpx, _ = oracle.build_codebook(px, ch, k, 3, 11)
    f, c = oracle.pixels_to_oklab(px, ch), oracle.pixels_to_oklab(cpx, ch)
    i_ref, e_ref = oracle.assign(f, c)
    i_gpu, e_gpu, st = enc.assign(f, c, assign_mode=2)
    assert bits_equal(i_gpu, i_ref) and bits_equal(e_gpu, e_ref)


def test_assign_tensor_ties_flat_and_nan(enc, oracle):
    rng = np.random.default_rng(5)
    c8 = oracle.pixels_to_oklab(rng.integers(0, 256, (8, 48), dtype=np.uint8), 3)
    cent = np.concatenate([c8] * 16)                       # every centroid 16 times: exact ties
    feat = np.concatenate([c8] * 1100)
    i_ref, e_ref = oracle.assign(feat, cent)
    i_gpu, e_gpu, _ = enc.assign(feat, cent, assign_mode=2)
    assert bits_equal(i_gpu, i_ref) and bits_equal(e_gpu, e_ref) and i_gpu.max() < 8
    flat_f, flat_c = np.tile(c8[:1], (9000, 1)), np.tile(c8[:1], (128, 1))
    i_gpu, e_gpu, _ = enc.assign(flat_f, flat_c, assign_mode=2)
    assert (i_gpu == 0).all() and (e_gpu == 0).all()
    # non-finite features: the reference's `d < best` never fires -> index 0, error FLT_MAX
    bad = feat[:9000].copy()
    bad[5, 3] = np.nan
    bad[77, 60] = np.inf
    i_x, e_x, _ = enc.assign(bad, cent, assign_mode=1)
    i_t, e_t, _ = enc.assign(bad, cent, assign_mode=2)
    assert bits_equal(i_t, i_x) and bits_equal(e_t, e_x) and i_t[5] == 0 and i_t[77] == 0


def test_assign_tensor_hdr_features(enc, oracle):
    """HDR Oklab features (large dynamic range) through the same engine."""
    img = synth.hdr_rgb(256, 128, seed=3)
    fl = enc.gather_tiles(img)
    f = oracle.floats_to_oklab(fl, 3)
    rng = np.random.default_rng(1)
    c = f[rng.choice(len(f), 1024, replace=False)]
    i_ref, e_ref = oracle.assign(f, c)
    i_gpu, e_gpu, _ = enc.assign(f, c, assign_mode=2)
    assert bits_equal(i_gpu, i_ref) and bits_equal(e_gpu, e_ref)


# ---- K2: seeding ------------------------------------------------------------------------------------------

@pytest.mark.parametrize("ch,n,k,seed", [(3, 65536, 512, 42), (4, 20000, 128, 7), (3, 1000, 64, 1), (1, 3000, 32, 9),
                                         (2, 3000, 32, 10), (3, 300, 1024, 3)])
def test_seed_ldr_vs_oracle(enc, oracle, ch, n, k, seed):
    rng = np.random.default_rng(seed)
    base = rng.integers(0, 256, (n // 8 + 1, 16 * ch), dtype=np.uint8)
    px = np.clip(base[rng.integers(0, len(base), n)].astype(np.int32) + rng.integers(-3, 4, (n, 16 * ch)), 0, 255).astype(np.uint8)
    s_ref, r_ref = oracle.seed_ldr(px, ch, k, seed)
    s_gpu, r_gpu = enc.seed_ldr(px, ch, k, seed)
    assert np.array_equal(s_gpu, s_ref) and r_gpu == r_ref


def test_seed_ldr_degenerate(enc, oracle):
    """Constant texture: the D^2 sum is 0 after round 1 -> every seed is seed 0 (reference :299-302);
    two-colour texture: the sum hits 0 after both colours are seeds."""
    px = np.tile(np.arange(48, dtype=np.uint8), (500, 1))
    s_ref, r_ref = oracle.seed_ldr(px, 3, 16, 5)
    s_gpu, r_gpu = enc.seed_ldr(px, 3, 16, 5)
    assert np.array_equal(s_gpu, s_ref) and r_gpu == r_ref == 0
    px[::3] = 200
    s_ref, r_ref = oracle.seed_ldr(px, 3, 16, 6)
    s_gpu, r_gpu = enc.seed_ldr(px, 3, 16, 6)
    assert np.array_equal(s_gpu, s_ref) and r_gpu == r_ref == 1


def test_seed_hdr_vs_oracle(enc, oracle):
    img = synth.hdr_rgb(256, 128, seed=8)
    feat = oracle.floats_to_oklab(synth.image_to_blocks(img), 3)
    for k, seed in [(64, 3), (128, 4)]:
        s_ref, r_ref = oracle.seed_hdr(feat, k, seed)
        s_gpu, r_gpu = enc.seed_hdr(feat, k, seed)
        # double sums are reduced in a different (fixed) tree than the reference's serial loop: the
        # chosen seeds can only differ if p lands within ~1e-16 relative of a bucket edge
        assert np.array_equal(s_gpu, s_ref) and r_gpu == r_ref


# ---- K4/K5: update ------------------------------------------------------------------------------------------

def test_update_centroids_exact_sums(enc, oracle):
    """c = float(exact sum) * (1.0f / count): checked against math.fsum (exactly rounded sums)."""
    import math
    rng = np.random.default_rng(11)
    n, k = 6000, 24
    feat = oracle.pixels_to_oklab(rng.integers(0, 256, (n, 64), dtype=np.uint8), 4)
    assign = rng.integers(0, k - 2, n).astype(np.uint32)       # clusters k-2, k-1 stay empty
    errors = rng.uniform(0, 1, n).astype(np.float32)
    errors[[17, 4000]] = 2.0                                    # tie: first maximum (17) wins, then 4000
    cent0 = np.zeros((k, 64), np.float32)
    cent, counts = enc.update_centroids(feat, assign, errors, cent0)
    assert np.array_equal(counts, np.bincount(assign, minlength=k).astype(np.uint32))
    for c in range(k - 2):
        rows = feat[assign == c].astype(np.float64)
        inv = np.float32(1.0) / np.float32(len(rows))
        want = np.array([np.float32(math.fsum(rows[:, j])) * inv for j in range(64)], np.float32)
        assert bits_equal(cent[c], want), c
    assert bits_equal(cent[k - 2], feat[17]) and bits_equal(cent[k - 1], feat[4000])


def test_update_centroids_sorted_path(enc, oracle):
    """n >= 8192 takes the sort-and-own accumulation kernel; sums stay exactly rounded, ragged tail chunk."""
    import math
    rng = np.random.default_rng(12)
    n, k = 8192 * 3 + 1234, 512
    feat = oracle.pixels_to_oklab(rng.integers(0, 256, (n, 48), dtype=np.uint8), 3)
    assign = np.minimum(rng.geometric(0.01, n) - 1, k - 3).astype(np.uint32)   # skewed sizes, k-2 and k-1 empty
    assign[: k - 2] = np.arange(k - 2)                                         # every other cluster non-empty
    errors = rng.uniform(0, 1, n).astype(np.float32)
    errors[[9000, 20000]] = 3.0
    cent, counts = enc.update_centroids(feat, assign, errors, np.zeros((k, 64), np.float32))
    assert np.array_equal(counts, np.bincount(assign, minlength=k).astype(np.uint32))
    for c in (0, 1, 7, 100, 300, k - 3):
        rows = feat[assign == c].astype(np.float64)
        if len(rows) == 0:
            continue
        inv = np.float32(1.0) / np.float32(len(rows))
        want = np.array([np.float32(math.fsum(rows[:, j])) * inv for j in range(64)], np.float32)
        assert bits_equal(cent[c], want), c
    assert bits_equal(cent[k - 2], feat[9000]) and bits_equal(cent[k - 1], feat[20000])


def test_gather_codebook(enc):
    rng = np.random.default_rng(2)
    for bs in (8, 16):
        cb = rng.integers(0, 256, (512, bs), dtype=np.uint8)
        idx = rng.integers(0, 512, 100001).astype(np.uint32)
        idx[5] = 70000  # out of range -> block left zero (reference skips it: src/vq_bcn_compressor.cpp:353-355)
        out = enc.gather_codebook(cb, idx)
        want = cb[np.minimum(idx, 511)]
        want[5] = 0
        assert bits_equal(out, want)


def test_assign_tensor_adversarial_near_ties(enc, oracle):
    """the margin's accumulation term is established by test.  Rows with large |x'||c'| whose two best centroids
    differ by a few ulps of the score, components of alternating sign and mixed magnitude (worst case for the fp32 accumulation
    inside the tensor core): the tensor engine must return the exact engine's (= the oracle's) indices and errors."""
    rng = np.random.default_rng(77)
    k, n = 512, 40000
    base = rng.uniform(-1.0, 1.0, (k // 2, 64)).astype(np.float32)
    base *= (1.0 + 6.0 * (rng.uniform(size=(k // 2, 64)) < 0.2)).astype(np.float32)          # a few large components
    base *= np.where(np.arange(64) % 2 == 0, 1.0, -1.0).astype(np.float32)                    # alternating signs
    twin = base.copy()
    j = rng.integers(0, 64, k // 2)
    twin[np.arange(k // 2), j] = np.nextafter(twin[np.arange(k // 2), j], np.float32(np.inf))  # one ulp apart in one column
    cent = np.concatenate([base, twin]).astype(np.float32)
    pick = rng.integers(0, k // 2, n)
    # rows close to the midpoint of a twin pair: both centroids are candidates, the exact re-rank must decide
    feat = (0.5 * (base[pick] + twin[pick]) + rng.normal(0, 1e-4, (n, 64))).astype(np.float32)
    i_e, e_e, _ = enc.assign(feat, cent, assign_mode=1)
    i_t, e_t, st = enc.assign(feat, cent, assign_mode=2)
    i_r, e_r = oracle.assign(feat, cent)
    assert np.array_equal(i_e, i_r) and bits_equal(e_e, e_r)
    assert np.array_equal(i_t, i_e) and bits_equal(e_t, e_e)
    assert st.rerank_blocks > n     # the twins really were ambiguous for the front-end
