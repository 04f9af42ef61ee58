"""TEST INFRASTRUCTURE — ctypes binding of the CPU restatement oracle/_build/libvqoracle.so
(oracle/vq_oracle.cpp).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module; the product never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "libvqoracle.so")

_u8p = C.POINTER(C.c_uint8)
_f32p = C.POINTER(C.c_float)
_u32p = C.POINTER(C.c_uint32)

METRIC_SAD, METRIC_OKLAB = 0, 1


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "_build/libvqoracle.so"])


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.orc_codebook_size.restype = C.c_uint32
        L.orc_codebook_size.argtypes = [C.c_float]
        L.orc_decode_bcn.restype = C.c_int
        L.orc_decode_bcn.argtypes = [_u8p, C.c_size_t, C.c_int, C.c_int, _u8p]
        L.orc_gather_tiles_f32.argtypes = [_f32p, C.c_uint32, C.c_uint32, C.c_uint32, _f32p]
        L.orc_pixels_to_oklab.argtypes = [_u8p, C.c_size_t, C.c_int, _f32p]
        L.orc_oklab_to_pixels.argtypes = [_f32p, C.c_size_t, C.c_int, _u8p]
        L.orc_floats_to_oklab.argtypes = [_f32p, C.c_size_t, C.c_int, _f32p]
        L.orc_oklab_to_floats.argtypes = [_f32p, C.c_size_t, C.c_int, _f32p]
        L.orc_oklab_distance.argtypes = [_f32p, C.c_size_t, _f32p, C.c_size_t, _f32p]
        L.orc_block_sad.argtypes = [_u8p, C.c_size_t, _u8p, C.c_size_t, C.c_int, _f32p]
        L.orc_assign.argtypes = [_f32p, C.c_size_t, _f32p, C.c_uint32, _u32p, _f32p]
        L.orc_seed_ldr.restype = C.c_uint32
        L.orc_seed_ldr.argtypes = [_u8p, C.c_size_t, C.c_int, C.c_uint32, C.c_uint32, _u32p]
        L.orc_seed_hdr.restype = C.c_uint32
        L.orc_seed_hdr.argtypes = [_f32p, C.c_size_t, C.c_uint32, C.c_uint32, _u32p]
        L.orc_build_codebook.restype = C.c_int
        L.orc_build_codebook.argtypes = [_u8p, C.c_size_t, C.c_int, C.c_uint32, C.c_uint32, C.c_int, C.c_float,
                                         C.c_uint32, C.c_int, _u8p, _u32p]
        L.orc_quantize_blocks.restype = C.c_int
        L.orc_quantize_blocks.argtypes = [_u8p, C.c_size_t, C.c_int, _u8p, C.c_uint32, C.c_int, _u32p]
        L.orc_build_codebook_hdr.restype = C.c_int
        L.orc_build_codebook_hdr.argtypes = [_f32p, C.c_size_t, C.c_int, C.c_uint32, C.c_uint32, C.c_float,
                                             C.c_uint32, C.c_int, _f32p, _u32p]
        L.orc_quantize_blocks_hdr.restype = C.c_int
        L.orc_quantize_blocks_hdr.argtypes = [_f32p, C.c_size_t, C.c_int, _f32p, C.c_uint32, _u32p]
        _lib = L
    return _lib


def _p(a, t):
    return a.ctypes.data_as(t)


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


BLOCK_BYTES = {1: 8, 2: 16, 3: 16, 4: 8, 5: 16, 6: 16, 7: 16}


def codebook_size(q):
    return int(lib().orc_codebook_size(q))


def decode_bcn(bc, fmt, ch):
    bc = _c(bc, np.uint8).reshape(-1)
    n = bc.size // BLOCK_BYTES[fmt]
    out = np.empty((n, 16 * ch), np.uint8)
    rc = lib().orc_decode_bcn(_p(bc, _u8p), n, fmt, ch, _p(out, _u8p))
    if rc:
        raise RuntimeError(f"orc_decode_bcn rc={rc}")
    return out


def gather_tiles_f32(img):
    img = _c(img, np.float32)
    h, w, ch = img.shape
    n = ((w + 3) // 4) * ((h + 3) // 4)
    out = np.empty((n, 16 * ch), np.float32)
    lib().orc_gather_tiles_f32(_p(img, _f32p), w, h, ch, _p(out, _f32p))
    return out


def pixels_to_oklab(px, ch):
    px = _c(px, np.uint8).reshape(-1, 16 * ch)
    out = np.empty((px.shape[0], 64), np.float32)
    lib().orc_pixels_to_oklab(_p(px, _u8p), px.shape[0], ch, _p(out, _f32p))
    return out


def oklab_to_pixels(ok, ch):
    ok = _c(ok, np.float32).reshape(-1, 64)
    out = np.empty((ok.shape[0], 16 * ch), np.uint8)
    lib().orc_oklab_to_pixels(_p(ok, _f32p), ok.shape[0], ch, _p(out, _u8p))
    return out


def floats_to_oklab(px, ch):
    px = _c(px, np.float32).reshape(-1, 16 * ch)
    out = np.empty((px.shape[0], 64), np.float32)
    lib().orc_floats_to_oklab(_p(px, _f32p), px.shape[0], ch, _p(out, _f32p))
    return out


def oklab_to_floats(ok, ch):
    ok = _c(ok, np.float32).reshape(-1, 64)
    out = np.empty((ok.shape[0], 16 * ch), np.float32)
    lib().orc_oklab_to_floats(_p(ok, _f32p), ok.shape[0], ch, _p(out, _f32p))
    return out


def oklab_distance(a, b):
    a = _c(a, np.float32).reshape(-1, 64)
    b = _c(b, np.float32).reshape(-1, 64)
    out = np.empty((a.shape[0], b.shape[0]), np.float32)
    lib().orc_oklab_distance(_p(a, _f32p), a.shape[0], _p(b, _f32p), b.shape[0], _p(out, _f32p))
    return out


def block_sad(a, b, ch):
    a = _c(a, np.uint8).reshape(-1, 16 * ch)
    b = _c(b, np.uint8).reshape(-1, 16 * ch)
    out = np.empty((a.shape[0], b.shape[0]), np.float32)
    lib().orc_block_sad(_p(a, _u8p), a.shape[0], _p(b, _u8p), b.shape[0], ch, _p(out, _f32p))
    return out


def assign(feat, cent):
    feat = _c(feat, np.float32).reshape(-1, 64)
    cent = _c(cent, np.float32).reshape(-1, 64)
    idx = np.empty(feat.shape[0], np.uint32)
    err = np.empty(feat.shape[0], np.float32)
    lib().orc_assign(_p(feat, _f32p), feat.shape[0], _p(cent, _f32p), cent.shape[0], _p(idx, _u32p), _p(err, _f32p))
    return idx, err


def seed_ldr(px, ch, k, seed):
    px = _c(px, np.uint8).reshape(-1, 16 * ch)
    out = np.empty(k, np.uint32)
    rounds = lib().orc_seed_ldr(_p(px, _u8p), px.shape[0], ch, k, seed, _p(out, _u32p))
    return out, int(rounds)


def seed_hdr(feat, k, seed):
    feat = _c(feat, np.float32).reshape(-1, 64)
    out = np.empty(k, np.uint32)
    rounds = lib().orc_seed_hdr(_p(feat, _f32p), feat.shape[0], k, seed, _p(out, _u32p))
    return out, int(rounds)


def build_codebook(px, ch, k, max_iter, seed, metric=METRIC_OKLAB, ratio=1.0, threads=1):
    px = _c(px, np.uint8).reshape(-1, 16 * ch)
    cent = np.zeros((k, 16 * ch), np.uint8)
    it = C.c_uint32(0)
    rc = lib().orc_build_codebook(_p(px, _u8p), px.shape[0], ch, k, max_iter, metric, ratio, seed, threads,
                                  _p(cent, _u8p), C.byref(it))
    if rc:
        raise RuntimeError(f"orc_build_codebook rc={rc}")
    return cent, int(it.value)


def quantize_blocks(px, ch, cent, metric=METRIC_OKLAB):
    px = _c(px, np.uint8).reshape(-1, 16 * ch)
    cent = _c(cent, np.uint8).reshape(-1, 16 * ch)
    idx = np.empty(px.shape[0], np.uint32)
    lib().orc_quantize_blocks(_p(px, _u8p), px.shape[0], ch, _p(cent, _u8p), cent.shape[0], metric, _p(idx, _u32p))
    return idx


def build_codebook_hdr(px, ch, k, max_iter, seed, ratio=1.0, threads=1):
    px = _c(px, np.float32).reshape(-1, 16 * ch)
    cent = np.zeros((k, 16 * ch), np.float32)
    it = C.c_uint32(0)
    lib().orc_build_codebook_hdr(_p(px, _f32p), px.shape[0], ch, k, max_iter, ratio, seed, threads,
                                 _p(cent, _f32p), C.byref(it))
    return cent, int(it.value)


def quantize_blocks_hdr(px, ch, cent):
    px = _c(px, np.float32).reshape(-1, 16 * ch)
    cent = _c(cent, np.float32).reshape(-1, 16 * ch)
    idx = np.empty(px.shape[0], np.uint32)
    lib().orc_quantize_blocks_hdr(_p(px, _f32p), px.shape[0], ch, _p(cent, _f32p), cent.shape[0], _p(idx, _u32p))
    return idx
