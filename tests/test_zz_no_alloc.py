"""yupvq_update_centroids_static: the centroid update for callers that may not allocate GPU memory.

Some users of this library run where cudaMalloc is off limits: inside an engine or tool whose allocator already owns
the whole device, in a frame loop that cannot take the process-wide stall of a driver allocation, or as a plugin that
is only ever handed buffers.  They cannot create a yupvq_ctx (yupvq_create allocates), so the library gives them one
context-free entry point whose working memory is a fixed 520 KiB array inside libyupvq.so.  These tests call it through
the C ABI with device buffers that the TEST allocates (torch), and require the same bits as the allocating entry point,
yupvq_update_centroids, on two different calls in a row -- the second would see the first call's sums if the shared array
were not cleared -- and once more on the first inputs.

The file name sorts last on purpose: `pytest -x` stops at the first failure, and a failure here must not hide the
parity tests.  It passes on its own in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import bits_equal

MAX_K = 1024   # YUPVQ_STATIC_MAX_CODEBOOK in include/yupvq.h


def _entry():
    from yuptexture_b200 import binding
    L = binding.lib()
    f = L.yupvq_update_centroids_static
    f.restype = C.c_int
    f.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p,
                  C.c_char_p, C.c_size_t]
    return f


def test_limit_in_header_matches_the_test():
    import os
    import re
    root = os.path.normpath(os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
    text = open(os.path.join(root, "include", "yupvq.h")).read()
    assert int(re.search(r"#define\s+YUPVQ_STATIC_MAX_CODEBOOK\s+(\d+)", text).group(1)) == MAX_K


def test_sizes_beyond_the_array_are_refused_before_any_gpu_work():
    """The size check comes first, so it needs no GPU: a codebook the array cannot hold is an error with a reason."""
    f = _entry()
    msg = C.create_string_buffer(256)
    dummy = C.c_void_p(256)   # never dereferenced: the call must stop at the size check
    assert f(0, None, dummy, 1000, dummy, dummy, MAX_K + 1, dummy, None, msg, len(msg)) != 0
    assert str(MAX_K).encode() in msg.value and b"codebook_size" in msg.value
    assert f(0, None, dummy, 0, dummy, dummy, 16, dummy, None, msg, len(msg)) != 0
    assert b"empty" in msg.value


def _case(seed, n, k, hot):
    """Random feature rows, assignments that leave clusters k-2 and k-1 empty, errors with two planted maxima."""
    rng = np.random.default_rng(seed)
    feat = rng.uniform(-0.4, 1.0, (n, 64)).astype(np.float32)
    assign = rng.integers(0, k - 2, n).astype(np.uint32)
    assign[: k - 2] = np.arange(k - 2)
    errors = rng.uniform(0, 1, n).astype(np.float32)
    errors[list(hot)] = 5.0   # equal maxima: the first one is taken first
    cent0 = rng.uniform(-1, 1, (k, 64)).astype(np.float32)
    return feat, assign, errors, cent0


def _run_static(f, torch, case, stream):
    feat, assign, errors, cent0 = case
    k = cent0.shape[0]
    d_feat = torch.from_numpy(feat).cuda()
    d_assign = torch.from_numpy(assign.view(np.int32)).cuda()
    d_err = torch.from_numpy(errors).cuda()
    d_cent = torch.from_numpy(cent0).cuda()
    d_counts = torch.full((k,), -1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()   # the uploads ran on torch's stream, the call uses `stream`
    msg = C.create_string_buffer(256)
    rc = f(0, C.c_void_p(stream.cuda_stream if stream is not None else 0), d_feat.data_ptr(), feat.shape[0], d_assign.data_ptr(),
           d_err.data_ptr(), k, d_cent.data_ptr(), d_counts.data_ptr(), msg, len(msg))
    assert rc == 0, msg.value
    return d_cent.cpu().numpy(), d_counts.cpu().numpy().view(np.uint32), d_err.cpu().numpy()


@pytest.mark.gpu
def test_static_update_equals_allocating_update_on_consecutive_calls():
    import torch
    from yuptexture_b200 import binding
    f = _entry()
    enc = binding.Encoder(0)
    big = _case(101, 8192 * 3 + 777, 512, (9000, 20000))      # sort-and-own accumulation kernel, ragged last chunk
    small = _case(102, 6000, 24, (17, 4000))                  # one-warp-per-block kernel; uses a prefix of the array
    side = torch.cuda.Stream()
    for case, hot, stream in ((big, (9000, 20000), None), (small, (17, 4000), side), (big, (9000, 20000), side)):
        feat, assign, errors, cent0 = case
        k = cent0.shape[0]
        want_cent, want_counts = enc.update_centroids(feat, assign, errors, cent0)
        cent, counts, err_after = _run_static(f, torch, case, stream)
        assert np.array_equal(counts, want_counts) and np.array_equal(counts, np.bincount(assign, minlength=k))
        assert bits_equal(cent, want_cent)                    # integer sums, same roundings: bit for bit
        assert bits_equal(cent[k - 2], feat[hot[0]]) and bits_equal(cent[k - 1], feat[hot[1]])
        want_err = errors.copy()
        want_err[list(hot)] = 0.0                             # the reseed zeroes the caller's errors in place
        assert bits_equal(err_after, want_err)
    enc.close()


@pytest.mark.gpu
def test_host_buffers_are_refused():
    """Staging a host buffer would take an allocation, so the entry point says no instead of letting a kernel fault."""
    import torch
    f = _entry()
    feat, assign, errors, cent0 = _case(103, 4096, 16, (1, 2))
    d_assign = torch.from_numpy(assign.view(np.int32)).cuda()
    d_err = torch.from_numpy(errors).cuda()
    d_cent = torch.from_numpy(cent0).cuda()
    torch.cuda.synchronize()
    msg = C.create_string_buffer(256)
    rc = f(0, None, feat.ctypes.data, feat.shape[0], d_assign.data_ptr(), d_err.data_ptr(), 16, d_cent.data_ptr(), None, msg, len(msg))
    assert rc != 0 and b"device" in msg.value
    assert bits_equal(d_cent.cpu().numpy(), cent0)            # nothing was written
