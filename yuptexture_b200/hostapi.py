"""ctypes binding of the plain-C entry points of libyuptexture.so (the C++ drop-in of VQBCnCompressor,
yuptexture_b200/host/yuptexture_host.cpp) for bench.py's whole-Compress breakdown and the tests.  Plumbing only."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libyuptexture.so")


class CompressionParams(C.Structure):
    """include/yuptexture.h CompressionParams == the reference's include/vq_bcn_types.h:110-126 (44 bytes)."""
    _fields_ = [("bcFormat", C.c_int32), ("bcQuality", C.c_float), ("zstdLevel", C.c_int32), ("numThreads", C.c_int32),
                ("alphaThreshold", C.c_uint8), ("useVQ", C.c_bool), ("useZstd", C.c_bool),
                ("vq_FastModeSampleRatio", C.c_float), ("quality", C.c_float), ("vq_Metric", C.c_int32),
                ("vq_min_cb_power", C.c_uint32), ("vq_max_cb_power", C.c_uint32), ("vq_maxIterations", C.c_uint32)]


assert C.sizeof(CompressionParams) == 44

_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_PATH)
        L.yuptexture_codec_name.restype = C.c_char_p
        L.yuptexture_load_codec.argtypes = [C.c_char_p]
        L.yuptexture_set_gpu_bcn_encode.argtypes = [C.c_int]
        L.yuptexture_last_encoder_name.restype = C.c_char_p
        L.yuptexture_compress_timed.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(CompressionParams),
                                                C.c_uint64, C.POINTER(C.c_double), C.POINTER(C.c_size_t), C.c_void_p, C.c_size_t,
                                                C.c_char_p, C.c_size_t]
        L.yuptexture_decompress_to_bcn.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t), C.c_void_p,
                                                   C.c_char_p, C.c_size_t]
        _lib = L
    return _lib


def codec_name():
    return lib().yuptexture_codec_name().decode()


def set_gpu_bcn_encode(mode):
    """-1: auto (GPU pre-pass only without a codec), 0: off, 1: on (BCnCompressor::SetGpuEncode)."""
    lib().yuptexture_set_gpu_bcn_encode(int(mode))


def last_encoder_name():
    return lib().yuptexture_last_encoder_name().decode()


def make_params(bc_format=7, quality=0.5, max_iterations=32, use_zstd=True, threads=16, bc_quality=1.0, metric=1, sample_ratio=1.0):
    p = CompressionParams(bc_format, bc_quality, 3, threads, 128, True, use_zstd, sample_ratio, quality, metric, 4, 10, max_iterations)
    return p


def compress_timed(img, bc_format=1, quality=0.8, max_iterations=32, use_zstd=True, threads=16, seed=42, bc_quality=1.0, metric=1,
                   want_stream=False):
    """VQBCnCompressor::Compress (uint8 image) or CompressHDR (float32 image) -> stage times in ms (+ the stream)."""
    is_hdr = img.dtype == np.float32
    img = np.ascontiguousarray(img)
    h, w, ch = img.shape
    p = make_params(bc_format, quality, max_iterations, use_zstd, threads, bc_quality, metric)
    ms = (C.c_double * 5)()
    size = C.c_size_t(0)
    cap = img.nbytes // 2 + (1 << 20) if want_stream else 0
    buf = np.empty(cap, np.uint8) if want_stream else None
    err = C.create_string_buffer(512)
    rc = lib().yuptexture_compress_timed(img.ctypes.data, 1 if is_hdr else 0, w, h, ch, C.byref(p), seed, ms, C.byref(size),
                                         buf.ctypes.data if want_stream else None, cap, err, 512)
    if rc:
        raise RuntimeError("yuptexture_compress_timed: " + err.value.decode())
    out = {"bcn_encode_ms": ms[0], "vq_core_ms": ms[1], "centroid_encodes_ms": ms[2], "zstd_ms": ms[3], "total_ms": ms[4],
           "stream_bytes": int(size.value)}
    return (out, buf[:size.value].copy()) if want_stream else out
