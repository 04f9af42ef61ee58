"""Test helper: the block encoders of yuptexture_b200/csrc/bcn_encode.cuh compiled as host C++ (tests/cpp/bcn_encode_host.cpp).
Test infrastructure only -- the product path is the CUDA kernel."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "cpp", "bcn_encode_host.cpp")
HDRS = [os.path.join(HERE, "..", "yuptexture_b200", "csrc", f) for f in ("bcn_encode.cuh", "bc1_single_tables.h")]
OUT = os.path.join(HERE, "cpp", "_build", "libbcnenc_host.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        newest = max(os.path.getmtime(p) for p in [SRC] + HDRS)
        if not os.path.exists(OUT) or os.path.getmtime(OUT) < newest:
            os.makedirs(os.path.dirname(OUT), exist_ok=True)
            # -ffp-contract=off: the kernel is built with --fmad=false, so both sides round every product and sum alike
            subprocess.check_call(["g++", "-O2", "-shared", "-fPIC", "-std=c++17", "-ffp-contract=off", SRC, "-o", OUT])
        _lib = C.CDLL(OUT)
        _lib.bcnenc_host_encode.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_uint32, C.c_void_p]
    return _lib


def encode(img, fmt, swap_rb=None, alpha_threshold=128):
    img = np.ascontiguousarray(img, np.uint8)
    h, w, ch = img.shape
    swap = (ch == 3) if swap_rb is None else bool(swap_rb)
    bs = 8 if fmt in (1, 4) else 16
    out = np.zeros((((w + 3) // 4) * ((h + 3) // 4), bs), np.uint8)
    rc = lib().bcnenc_host_encode(img.ctypes.data_as(C.c_void_p), w, h, ch, fmt, int(swap), alpha_threshold, out.ctypes.data_as(C.c_void_p))
    if rc:
        raise RuntimeError("bcnenc_host_encode failed")
    return out
