"""TEST INFRASTRUCTURE — ctypes binding of oracle/_ref/libyupref.so (the unmodified reference
VQ path built by oracle/build_ref.sh).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module; the product never does.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libyupref.so")

_u8p = C.POINTER(C.c_uint8)
_f32p = C.POINTER(C.c_float)
_u32p = C.POINTER(C.c_uint32)
_f64p = C.POINTER(C.c_double)

BC = {"BC1": 1, "BC2": 2, "BC3": 3, "BC4": 4, "BC5": 5, "BC6H": 6, "BC7": 7}
BLOCK_BYTES = {1: 8, 2: 16, 3: 16, 4: 8, 5: 16, 6: 16, 7: 16}
METRIC_SAD, METRIC_OKLAB = 0, 1


def available():
    return os.path.exists(LIB_PATH)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError(f"{LIB_PATH} missing: run oracle/build_ref.sh where /root/reference exists")
        _lib = C.CDLL(LIB_PATH)
        _lib.ref_codebook_size.restype = C.c_uint32
        _lib.ref_codebook_size.argtypes = [C.c_float]
        _lib.ref_bcn_encode.restype = C.c_size_t
        _lib.ref_bcn_encode.argtypes = [_u8p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_float, C.c_int,
                                        C.c_int, _u8p, C.c_size_t]
        _lib.ref_bcn_encode_hdr.restype = C.c_size_t
        _lib.ref_bcn_encode_hdr.argtypes = [_f32p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_float,
                                            C.c_int, _u8p, C.c_size_t]
        _lib.ref_bcn_decode_blocks.restype = C.c_int
        _lib.ref_bcn_decode_blocks.argtypes = [_u8p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, _u8p]
        _lib.ref_bcn_decode_each.restype = C.c_int
        _lib.ref_bcn_decode_each.argtypes = [_u8p, C.c_size_t, C.c_uint32, C.c_int, _u8p]
        _lib.ref_pixels_to_oklab.argtypes = [_u8p, C.c_size_t, C.c_uint32, _f32p]
        _lib.ref_oklab_to_pixels.argtypes = [_f32p, C.c_size_t, C.c_uint32, _u8p]
        _lib.ref_floats_to_oklab.argtypes = [_f32p, C.c_size_t, C.c_uint32, _f32p]
        _lib.ref_oklab_to_floats.argtypes = [_f32p, C.c_size_t, C.c_uint32, _f32p]
        _lib.ref_oklab_distance.argtypes = [_f32p, C.c_size_t, _f32p, C.c_size_t, _f32p]
        _lib.ref_block_sad.argtypes = [_u8p, C.c_size_t, _u8p, C.c_size_t, C.c_uint32, _f32p]
        _lib.ref_build_codebook.restype = C.c_int
        _lib.ref_build_codebook.argtypes = [_u8p, C.c_size_t, C.c_uint32, C.c_int, C.c_float, C.c_uint32, C.c_int,
                                            C.c_float, C.c_uint32, C.c_int, C.c_float, _u8p, _u8p, _f64p]
        _lib.ref_quantize_blocks.restype = C.c_int
        _lib.ref_quantize_blocks.argtypes = [_u8p, C.c_size_t, C.c_uint32, _u8p, C.c_uint32, C.c_int, C.c_int,
                                             _u32p, _f64p]
        _lib.ref_build_codebook_hdr.restype = C.c_int
        _lib.ref_build_codebook_hdr.argtypes = [_f32p, C.c_size_t, C.c_uint32, C.c_int, C.c_float, C.c_uint32,
                                                C.c_float, C.c_uint32, C.c_int, C.c_float, _f32p, _u8p, _f64p]
        _lib.ref_quantize_blocks_hdr.restype = C.c_int
        _lib.ref_quantize_blocks_hdr.argtypes = [_f32p, C.c_size_t, C.c_uint32, _f32p, C.c_uint32, C.c_int,
                                                 _u32p, _f64p]
        _lib.ref_encode_single_block.restype = C.c_size_t
        _lib.ref_encode_single_block.argtypes = [_u8p, C.c_uint32, C.c_int, C.c_float, C.c_int, _u8p]
        _lib.ref_num_threads.restype = C.c_int
    return _lib


def _p(a, t):
    return a.ctypes.data_as(t)


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def codebook_size(quality):
    return int(lib().ref_codebook_size(quality))


def bcn_encode(img, fmt, quality=1.0, threads=8, alpha_threshold=128):
    """img: (H, W, ch) uint8 -> BCn bytes (np.uint8, n_blocks*blockSize)."""
    img = _c(img, np.uint8)
    h, w, ch = img.shape
    n = ((w + 3) // 4) * ((h + 3) // 4) * BLOCK_BYTES[fmt]
    out = np.empty(n, np.uint8)
    got = lib().ref_bcn_encode(_p(img, _u8p), w, h, ch, fmt, quality, threads, alpha_threshold, _p(out, _u8p), n)
    if got != n:
        raise RuntimeError("reference BCn encode failed")
    return out


def bcn_decode_blocks(bc, w, h, ch, fmt):
    """-> (n_blocks, 16*ch) uint8 pixel blocks, block order row-major."""
    bc = _c(bc, np.uint8)
    n = ((w + 3) // 4) * ((h + 3) // 4)
    out = np.empty((n, 16 * ch), np.uint8)
    if lib().ref_bcn_decode_blocks(_p(bc, _u8p), w, h, ch, fmt, _p(out, _u8p)):
        raise RuntimeError("reference BCn decode failed")
    return out


def bcn_decode_each(bc, ch, fmt):
    bc = _c(bc, np.uint8)
    n = bc.size // BLOCK_BYTES[fmt]
    out = np.empty((n, 16 * ch), np.uint8)
    if lib().ref_bcn_decode_each(_p(bc, _u8p), n, ch, fmt, _p(out, _u8p)):
        raise RuntimeError("reference per-block BCn decode failed")
    return out


def pixels_to_oklab(blocks, ch):
    blocks = _c(blocks, np.uint8).reshape(-1, 16 * ch)
    out = np.empty((blocks.shape[0], 64), np.float32)
    lib().ref_pixels_to_oklab(_p(blocks, _u8p), blocks.shape[0], ch, _p(out, _f32p))
    return out


def oklab_to_pixels(oklab, ch):
    oklab = _c(oklab, np.float32).reshape(-1, 64)
    out = np.empty((oklab.shape[0], 16 * ch), np.uint8)
    lib().ref_oklab_to_pixels(_p(oklab, _f32p), oklab.shape[0], ch, _p(out, _u8p))
    return out


def floats_to_oklab(blocks, ch):
    blocks = _c(blocks, np.float32).reshape(-1, 16 * ch)
    out = np.empty((blocks.shape[0], 64), np.float32)
    lib().ref_floats_to_oklab(_p(blocks, _f32p), blocks.shape[0], ch, _p(out, _f32p))
    return out


def oklab_to_floats(oklab, ch):
    oklab = _c(oklab, np.float32).reshape(-1, 64)
    out = np.empty((oklab.shape[0], 16 * ch), np.float32)
    lib().ref_oklab_to_floats(_p(oklab, _f32p), oklab.shape[0], ch, _p(out, _f32p))
    return out


def oklab_distance(a, b):
    a = _c(a, np.float32).reshape(-1, 64)
    b = _c(b, np.float32).reshape(-1, 64)
    out = np.empty((a.shape[0], b.shape[0]), np.float32)
    lib().ref_oklab_distance(_p(a, _f32p), a.shape[0], _p(b, _f32p), b.shape[0], _p(out, _f32p))
    return out


def block_sad(a, b, ch):
    a = _c(a, np.uint8).reshape(-1, 16 * ch)
    b = _c(b, np.uint8).reshape(-1, 16 * ch)
    out = np.empty((a.shape[0], b.shape[0]), np.float32)
    lib().ref_block_sad(_p(a, _u8p), a.shape[0], _p(b, _u8p), b.shape[0], ch, _p(out, _f32p))
    return out


def build_codebook(blocks, ch, fmt, quality, max_iter, seed, metric=METRIC_OKLAB, ratio=1.0, threads=8,
                   bc_quality=1.0, want_entries=False):
    """-> (pixel_centroids (K,16*ch) u8, entries or None, seconds)."""
    blocks = _c(blocks, np.uint8).reshape(-1, 16 * ch)
    k = codebook_size(quality)
    cent = np.zeros((k, 16 * ch), np.uint8)
    ent = np.zeros((k, BLOCK_BYTES[fmt]), np.uint8) if want_entries else None
    t = np.zeros(1, np.float64)
    got = lib().ref_build_codebook(_p(blocks, _u8p), blocks.shape[0], ch, fmt, quality, max_iter, metric, ratio,
                                   seed, threads, bc_quality, _p(cent, _u8p),
                                   _p(ent, _u8p) if want_entries else None, _p(t, _f64p))
    assert got == k
    return cent, ent, float(t[0])


def quantize_blocks(blocks, ch, centroids, metric=METRIC_OKLAB, threads=8):
    blocks = _c(blocks, np.uint8).reshape(-1, 16 * ch)
    centroids = _c(centroids, np.uint8).reshape(-1, 16 * ch)
    idx = np.empty(blocks.shape[0], np.uint32)
    t = np.zeros(1, np.float64)
    lib().ref_quantize_blocks(_p(blocks, _u8p), blocks.shape[0], ch, _p(centroids, _u8p), centroids.shape[0],
                              metric, threads, _p(idx, _u32p), _p(t, _f64p))
    return idx, float(t[0])


def build_codebook_hdr(blocks, ch, quality, max_iter, seed, ratio=1.0, threads=8, bc_quality=0.25,
                       want_entries=False):
    blocks = _c(blocks, np.float32).reshape(-1, 16 * ch)
    k = codebook_size(quality)
    cent = np.zeros((k, 16 * ch), np.float32)
    ent = np.zeros((k, 16), np.uint8) if want_entries else None
    t = np.zeros(1, np.float64)
    got = lib().ref_build_codebook_hdr(_p(blocks, _f32p), blocks.shape[0], ch, BC["BC6H"], quality, max_iter,
                                       ratio, seed, threads, bc_quality, _p(cent, _f32p),
                                       _p(ent, _u8p) if want_entries else None, _p(t, _f64p))
    assert got == k
    return cent, ent, float(t[0])


def quantize_blocks_hdr(blocks, ch, centroids, threads=8):
    blocks = _c(blocks, np.float32).reshape(-1, 16 * ch)
    centroids = _c(centroids, np.float32).reshape(-1, 16 * ch)
    idx = np.empty(blocks.shape[0], np.uint32)
    t = np.zeros(1, np.float64)
    lib().ref_quantize_blocks_hdr(_p(blocks, _f32p), blocks.shape[0], ch, _p(centroids, _f32p),
                                  centroids.shape[0], threads, _p(idx, _u32p), _p(t, _f64p))
    return idx, float(t[0])


def encode_blocks(pixel_blocks, ch, fmt, bc_quality=1.0, threads=8):
    """K calls of the reference's 4x4 encode (CompressSingleBlock, src/vq_encoder.cpp:248-250) -> (K, blockSize) u8."""
    pixel_blocks = _c(pixel_blocks, np.uint8).reshape(-1, 16 * ch)
    out = np.zeros((pixel_blocks.shape[0], BLOCK_BYTES[fmt]), np.uint8)
    for k in range(pixel_blocks.shape[0]):
        lib().ref_encode_single_block(_p(pixel_blocks[k], _u8p), ch, fmt, bc_quality, threads, _p(out[k], _u8p))
    return out


def num_threads():
    return int(lib().ref_num_threads())
