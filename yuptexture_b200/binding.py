"""ctypes binding of the C-ABI library libyupvq.so (include/yupvq.h) for the Python test and
bench drivers.  The product is the library; this module is plumbing.  It fails loudly when the
CUDA library is missing — there is no CPU path.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("YUPVQ_LIB") or os.path.join(_HERE, "libyupvq.so")

BC = {"BC1": 1, "BC2": 2, "BC3": 3, "BC4": 4, "BC5": 5, "BC6H": 6, "BC7": 7}
BLOCK_BYTES = {1: 8, 2: 16, 3: 16, 4: 8, 5: 16, 6: 16, 7: 16}
METRIC_SAD, METRIC_OKLAB = 0, 1
ASSIGN_AUTO, ASSIGN_EXACT, ASSIGN_TENSOR = 0, 1, 2


class Config(C.Structure):
    _fields_ = [("codebook_size", C.c_uint32), ("max_iterations", C.c_uint32), ("metric", C.c_int32),
                ("sample_ratio", C.c_float), ("seed", C.c_uint64), ("assign_mode", C.c_int32),
                ("reserved", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("n_blocks", C.c_uint32), ("codebook_size", C.c_uint32), ("lloyd_passes", C.c_uint32),
                ("seed_rounds", C.c_uint32), ("reseeds", C.c_uint32), ("rerank_blocks", C.c_uint32),
                ("ms_h2d", C.c_float), ("ms_decode", C.c_float), ("ms_seed", C.c_float), ("ms_lloyd", C.c_float),
                ("ms_finalize", C.c_float), ("ms_quantize", C.c_float), ("ms_d2h", C.c_float),
                ("ms_total", C.c_float), ("gpu_launches", C.c_uint32), ("rerank_overflow", C.c_uint32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if k != "reserved"}


_vp = C.c_void_p
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `make -C yuptexture_b200/csrc` "
                               "(python -c 'import __graft_entry__ as g; g.build()'); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        L.yupvq_create.argtypes = [C.c_int, C.POINTER(_vp)]
        L.yupvq_destroy.argtypes = [_vp]
        L.yupvq_last_error.restype = C.c_char_p
        L.yupvq_last_error.argtypes = [_vp]
        L.yupvq_version.restype = C.c_char_p
        L.yupvq_timer_start.argtypes = [_vp]
        L.yupvq_timer_stop.argtypes = [_vp, C.POINTER(C.c_float)]
        L.yupvq_codebook_size.restype = C.c_uint32
        L.yupvq_codebook_size.argtypes = [C.c_float]
        L.yupvq_config_init.argtypes = [C.POINTER(Config), C.c_float, C.c_uint32, C.c_int, C.c_float, C.c_uint64]
        cfgp, stp = C.POINTER(Config), C.POINTER(Stats)
        L.yupvq_encode_bcn.argtypes = [_vp, _vp, C.c_size_t, C.c_int, C.c_int, cfgp, _vp, _vp, stp]
        L.yupvq_encode_hdr.argtypes = [_vp, _vp, C.c_uint32, C.c_uint32, C.c_int, cfgp, _vp, _vp, stp]
        L.yupvq_encode_bcn_batch.argtypes = [_vp, C.POINTER(_vp), C.c_size_t, C.c_size_t, C.c_int, C.c_int, cfgp,
                                             C.POINTER(_vp), C.POINTER(_vp), stp]
        L.yupvq_build_codebook.argtypes = [_vp, _vp, C.c_size_t, C.c_int, cfgp, _vp, stp]
        L.yupvq_quantize_blocks.argtypes = [_vp, _vp, C.c_size_t, C.c_int, _vp, C.c_uint32, C.c_int, C.c_int, _vp, stp]
        L.yupvq_build_codebook_hdr.argtypes = [_vp, _vp, C.c_size_t, C.c_int, cfgp, _vp, stp]
        L.yupvq_quantize_blocks_hdr.argtypes = [_vp, _vp, C.c_size_t, C.c_int, _vp, C.c_uint32, C.c_int, _vp, stp]
        L.yupvq_decode_bcn.argtypes = [_vp, _vp, C.c_size_t, C.c_int, C.c_int, _vp]
        L.yupvq_pixels_to_oklab.argtypes = [_vp, _vp, C.c_size_t, C.c_int, _vp]
        L.yupvq_oklab_to_pixels.argtypes = [_vp, _vp, C.c_size_t, C.c_int, _vp]
        L.yupvq_floats_to_oklab.argtypes = [_vp, _vp, C.c_size_t, C.c_int, _vp]
        L.yupvq_oklab_to_floats.argtypes = [_vp, _vp, C.c_size_t, C.c_int, _vp]
        L.yupvq_gather_tiles.argtypes = [_vp, _vp, C.c_uint32, C.c_uint32, C.c_int, _vp]
        L.yupvq_assign.argtypes = [_vp, _vp, C.c_size_t, _vp, C.c_uint32, C.c_int, _vp, _vp, stp]
        L.yupvq_seed_ldr.argtypes = [_vp, _vp, C.c_size_t, C.c_int, C.c_uint32, C.c_uint64, _vp, C.POINTER(C.c_uint32)]
        L.yupvq_seed_hdr.argtypes = [_vp, _vp, C.c_size_t, C.c_uint32, C.c_uint64, _vp, C.POINTER(C.c_uint32)]
        L.yupvq_update_centroids.argtypes = [_vp, _vp, C.c_size_t, _vp, _vp, C.c_uint32, _vp, _vp]
        L.yupvq_gather_codebook.argtypes = [_vp, _vp, C.c_uint32, C.c_uint32, _vp, C.c_size_t, _vp]
        L.yupvq_bcn_encode_supported.argtypes = [C.c_int]
        L.yupvq_bcn_encode_image.argtypes = [_vp, _vp, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_int, _vp, C.POINTER(C.c_float)]
        L.yupvq_compress_image.argtypes = [_vp, _vp, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_int, cfgp, _vp, _vp, _vp, stp,
                                           C.POINTER(C.c_float)]
        L.yupvq_profile.argtypes = [_vp, C.c_int]
        L.yupvq_profile_read.argtypes = [_vp, C.POINTER(C.c_float), C.POINTER(C.c_uint32)]
        L.yupvq_profile_stages.argtypes = [_vp, C.POINTER(C.c_float), C.POINTER(C.c_uint32)]
        L.yupvq_exchange_path.argtypes = [_vp]
        L.yupvq_comm_unique_id.argtypes = [_vp, _vp]
        L.yupvq_comm_init.argtypes = [_vp, _vp, C.c_int, C.c_int]
        L.yupvq_encode_bcn_sharded.argtypes = [_vp, _vp, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.c_int, cfgp,
                                               _vp, _vp, stp]
        L.yupvq_encode_bcn_batch_sharded.argtypes = [_vp, C.POINTER(_vp), C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t,
                                                     C.c_int, C.c_int, cfgp, C.POINTER(_vp), C.POINTER(_vp), stp, _vp]
        L.yupvq_encode_hdr_batch_sharded.argtypes = [_vp, C.POINTER(_vp), C.c_size_t, C.c_uint32, C.c_uint32, C.c_int, C.c_size_t,
                                                     C.c_size_t, cfgp, C.POINTER(_vp), C.POINTER(_vp), stp]
        L.yupvq_encode_hdr_sharded.argtypes = [_vp, _vp, C.c_uint32, C.c_uint32, C.c_int, C.c_size_t, C.c_size_t, cfgp, _vp, _vp, stp]
        _lib = L
    return _lib


def codebook_size(quality):
    return int(lib().yupvq_codebook_size(quality))


def make_config(quality=None, codebook=None, max_iterations=32, metric=METRIC_OKLAB, sample_ratio=1.0, seed=0,
                assign_mode=ASSIGN_AUTO):
    cfg = Config()
    lib().yupvq_config_init(C.byref(cfg), 0.5 if quality is None else quality, max_iterations, metric, sample_ratio, seed)
    if codebook is not None:
        cfg.codebook_size = codebook
    cfg.assign_mode = assign_mode
    return cfg


def _ptr(a):
    """numpy array -> host pointer; int -> raw (device) pointer; None -> NULL."""
    if a is None:
        return None
    if isinstance(a, int):
        return _vp(a)
    return _vp(a.ctypes.data)


def _c(a, dt):
    if isinstance(a, int) or a is None:
        return a
    return np.ascontiguousarray(a, dtype=dt)


class Encoder:
    """One yupvq_ctx bound to one CUDA device."""

    def __init__(self, device=0):
        self._L = lib()
        self._ctx = _vp()
        if self._L.yupvq_create(device, C.byref(self._ctx)):
            raise RuntimeError("yupvq_create: " + self._L.yupvq_last_error(None).decode())
        self.device = device

    def close(self):
        if self._ctx:
            self._L.yupvq_destroy(self._ctx)
            self._ctx = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc, what):
        if rc:
            raise RuntimeError(f"{what}: {self._L.yupvq_last_error(self._ctx).decode()}")

    def timer_start(self):
        self._ck(self._L.yupvq_timer_start(self._ctx), "yupvq_timer_start")

    def timer_stop(self):
        ms = C.c_float(0)
        self._ck(self._L.yupvq_timer_stop(self._ctx, C.byref(ms)), "yupvq_timer_stop")
        return float(ms.value)

    # ---- whole path
    def encode_bcn(self, bc, n_blocks, fmt, ch, cfg, out_centroids=None, out_indices=None):
        """bc / outputs: numpy arrays (host) or int device pointers.  Returns (centroids, indices, stats)."""
        bc = _c(bc, np.uint8)
        k = cfg.codebook_size
        if out_centroids is None:
            out_centroids = np.empty((k, 16 * ch), np.uint8)
        if out_indices is None:
            out_indices = np.empty(n_blocks, np.uint32)
        st = Stats()
        self._ck(self._L.yupvq_encode_bcn(self._ctx, _ptr(bc), n_blocks, fmt, ch, C.byref(cfg), _ptr(out_centroids),
                                          _ptr(out_indices), C.byref(st)), "yupvq_encode_bcn")
        return out_centroids, out_indices, st

    def encode_hdr(self, image, width, height, ch, cfg, out_centroids=None, out_indices=None):
        image = _c(image, np.float32)
        k = cfg.codebook_size
        n = ((width + 3) // 4) * ((height + 3) // 4)
        if out_centroids is None:
            out_centroids = np.empty((k, 16 * ch), np.float32)
        if out_indices is None:
            out_indices = np.empty(n, np.uint32)
        st = Stats()
        self._ck(self._L.yupvq_encode_hdr(self._ctx, _ptr(image), width, height, ch, C.byref(cfg), _ptr(out_centroids),
                                          _ptr(out_indices), C.byref(st)), "yupvq_encode_hdr")
        return out_centroids, out_indices, st

    def encode_bcn_sharded(self, bc, n_total, first, n_local, fmt, ch, cfg, out_centroids=None, out_indices=None):
        bc = _c(bc, np.uint8)
        if out_centroids is None:
            out_centroids = np.empty((cfg.codebook_size, 16 * ch), np.uint8)
        if out_indices is None:
            out_indices = np.empty(n_local, np.uint32)
        st = Stats()
        self._ck(self._L.yupvq_encode_bcn_sharded(self._ctx, _ptr(bc), n_total, first, n_local, fmt, ch, C.byref(cfg),
                                                  _ptr(out_centroids), _ptr(out_indices), C.byref(st)),
                 "yupvq_encode_bcn_sharded")
        return out_centroids, out_indices, st

    def exchange_path(self):
        """How the last *_sharded call exchanged its partial sums: 'none', 'peer_one_shot', 'peer_two_shot' or 'nccl'."""
        return ("none", "peer_one_shot", "peer_two_shot", "nccl")[self._L.yupvq_exchange_path(self._ctx)]

    def profile(self, enable):
        self._ck(self._L.yupvq_profile(self._ctx, 1 if enable else 0), "yupvq_profile")

    def profile_stages(self):
        """-> ({front_end, rerank, update, tail: ms summed over the profiled passes}, passes)"""
        ms = (C.c_float * 4)()
        n = C.c_uint32(0)
        self._ck(self._L.yupvq_profile_stages(self._ctx, ms, C.byref(n)), "yupvq_profile_stages")
        return {"front_end": ms[0], "rerank": ms[1], "update": ms[2], "tail_and_exchange": ms[3]}, n.value

    def profile_read(self):
        """(summed ms of the tensor front-end launches since the last read, number of launches)"""
        ms, n = C.c_float(0), C.c_uint32(0)
        self._ck(self._L.yupvq_profile_read(self._ctx, C.byref(ms), C.byref(n)), "yupvq_profile_read")
        return ms.value, n.value

    def encode_bcn_batch_sharded(self, textures, n_total, first, n_local, fmt, ch, cfg, out_centroids=None, out_indices=None,
                                 shard_only=None):
        """Lock-step batch (config 5): `textures` = list of whole-texture BCn arrays or raw device pointers
        (shard_only[t] true: just this rank's blocks [first, first + n_local) of texture t);
        outputs are lists (numpy arrays or raw device pointers), one per texture."""
        t_n = len(textures)
        textures = [x if isinstance(x, int) else _c(x, np.uint8) for x in textures]
        if out_centroids is None:
            out_centroids = [np.empty((cfg.codebook_size, 16 * ch), np.uint8) for _ in range(t_n)]
        if out_indices is None:
            out_indices = [np.empty(n_local, np.uint32) for _ in range(t_n)]
        arr = lambda xs: (_vp * t_n)(*[None if x is None else _ptr(x).value for x in xs])
        st = Stats()
        flags = None if shard_only is None else np.ascontiguousarray([1 if x else 0 for x in shard_only], dtype=np.uint8)
        self._ck(self._L.yupvq_encode_bcn_batch_sharded(self._ctx, arr(textures), t_n, n_total, first, n_local, fmt, ch,
                                                        C.byref(cfg), arr(out_centroids), arr(out_indices), C.byref(st),
                                                        _ptr(flags)),
                 "yupvq_encode_bcn_batch_sharded")
        return out_centroids, out_indices, st

    def encode_hdr_batch_sharded(self, images, width, height, ch, first, n_local, cfg, out_centroids=None, out_indices=None):
        """HDR images over the ranks (config 4): every rank passes every image whole (numpy arrays or raw device pointers)."""
        t_n = len(images)
        images = [x if isinstance(x, int) else _c(x, np.float32) for x in images]
        if out_centroids is None:
            out_centroids = [np.empty((cfg.codebook_size, 16 * ch), np.float32) for _ in range(t_n)]
        if out_indices is None:
            out_indices = [np.empty(n_local, np.uint32) for _ in range(t_n)]
        arr = lambda xs: (_vp * t_n)(*[None if x is None else _ptr(x).value for x in xs])
        st = Stats()
        self._ck(self._L.yupvq_encode_hdr_batch_sharded(self._ctx, arr(images), t_n, width, height, ch, first, n_local, C.byref(cfg),
                                                        arr(out_centroids), arr(out_indices), C.byref(st)),
                 "yupvq_encode_hdr_batch_sharded")
        return out_centroids, out_indices, st

    def encode_hdr_sharded(self, image, width, height, ch, first, n_local, cfg, out_centroids=None, out_indices=None):
        image = _c(image, np.float32)
        if out_centroids is None:
            out_centroids = np.empty((cfg.codebook_size, 16 * ch), np.float32)
        if out_indices is None:
            out_indices = np.empty(n_local, np.uint32)
        st = Stats()
        self._ck(self._L.yupvq_encode_hdr_sharded(self._ctx, _ptr(image), width, height, ch, first, n_local, C.byref(cfg),
                                                  _ptr(out_centroids), _ptr(out_indices), C.byref(st)), "yupvq_encode_hdr_sharded")
        return out_centroids, out_indices, st

    # ---- the four VQEncoder calls
    def build_codebook(self, px, ch, cfg):
        px = _c(px, np.uint8).reshape(-1, 16 * ch)
        out = np.empty((cfg.codebook_size, 16 * ch), np.uint8)
        st = Stats()
        self._ck(self._L.yupvq_build_codebook(self._ctx, _ptr(px), px.shape[0], ch, C.byref(cfg), _ptr(out), C.byref(st)),
                 "yupvq_build_codebook")
        return out, st

    def quantize_blocks(self, px, ch, cent, metric=METRIC_OKLAB, assign_mode=ASSIGN_AUTO):
        px = _c(px, np.uint8).reshape(-1, 16 * ch)
        cent = _c(cent, np.uint8).reshape(-1, 16 * ch)
        out = np.empty(px.shape[0], np.uint32)
        st = Stats()
        self._ck(self._L.yupvq_quantize_blocks(self._ctx, _ptr(px), px.shape[0], ch, _ptr(cent), cent.shape[0], metric,
                                               assign_mode, _ptr(out), C.byref(st)), "yupvq_quantize_blocks")
        return out, st

    def build_codebook_hdr(self, px, ch, cfg):
        px = _c(px, np.float32).reshape(-1, 16 * ch)
        out = np.empty((cfg.codebook_size, 16 * ch), np.float32)
        st = Stats()
        self._ck(self._L.yupvq_build_codebook_hdr(self._ctx, _ptr(px), px.shape[0], ch, C.byref(cfg), _ptr(out),
                                                  C.byref(st)), "yupvq_build_codebook_hdr")
        return out, st

    def quantize_blocks_hdr(self, px, ch, cent, assign_mode=ASSIGN_AUTO):
        px = _c(px, np.float32).reshape(-1, 16 * ch)
        cent = _c(cent, np.float32).reshape(-1, 16 * ch)
        out = np.empty(px.shape[0], np.uint32)
        st = Stats()
        self._ck(self._L.yupvq_quantize_blocks_hdr(self._ctx, _ptr(px), px.shape[0], ch, _ptr(cent), cent.shape[0],
                                                   assign_mode, _ptr(out), C.byref(st)), "yupvq_quantize_blocks_hdr")
        return out, st

    # ---- kernel level
    def decode_bcn(self, bc, fmt, ch):
        bc = _c(bc, np.uint8).reshape(-1)
        n = bc.size // BLOCK_BYTES[fmt]
        out = np.empty((n, 16 * ch), np.uint8)
        self._ck(self._L.yupvq_decode_bcn(self._ctx, _ptr(bc), n, fmt, ch, _ptr(out)), "yupvq_decode_bcn")
        return out

    def pixels_to_oklab(self, px, ch):
        px = _c(px, np.uint8).reshape(-1, 16 * ch)
        out = np.empty((px.shape[0], 64), np.float32)
        self._ck(self._L.yupvq_pixels_to_oklab(self._ctx, _ptr(px), px.shape[0], ch, _ptr(out)), "yupvq_pixels_to_oklab")
        return out

    def oklab_to_pixels(self, ok, ch):
        ok = _c(ok, np.float32).reshape(-1, 64)
        out = np.empty((ok.shape[0], 16 * ch), np.uint8)
        self._ck(self._L.yupvq_oklab_to_pixels(self._ctx, _ptr(ok), ok.shape[0], ch, _ptr(out)), "yupvq_oklab_to_pixels")
        return out

    def floats_to_oklab(self, px, ch):
        px = _c(px, np.float32).reshape(-1, 16 * ch)
        out = np.empty((px.shape[0], 64), np.float32)
        self._ck(self._L.yupvq_floats_to_oklab(self._ctx, _ptr(px), px.shape[0], ch, _ptr(out)), "yupvq_floats_to_oklab")
        return out

    def oklab_to_floats(self, ok, ch):
        ok = _c(ok, np.float32).reshape(-1, 64)
        out = np.empty((ok.shape[0], 16 * ch), np.float32)
        self._ck(self._L.yupvq_oklab_to_floats(self._ctx, _ptr(ok), ok.shape[0], ch, _ptr(out)), "yupvq_oklab_to_floats")
        return out

    def gather_tiles(self, img):
        img = _c(img, np.float32)
        h, w, ch = img.shape
        n = ((w + 3) // 4) * ((h + 3) // 4)
        out = np.empty((n, 16 * ch), np.float32)
        self._ck(self._L.yupvq_gather_tiles(self._ctx, _ptr(img), w, h, ch, _ptr(out)), "yupvq_gather_tiles")
        return out

    def assign(self, feat, cent, assign_mode=ASSIGN_AUTO):
        feat = _c(feat, np.float32).reshape(-1, 64)
        cent = _c(cent, np.float32).reshape(-1, 64)
        idx = np.empty(feat.shape[0], np.uint32)
        err = np.empty(feat.shape[0], np.float32)
        st = Stats()
        self._ck(self._L.yupvq_assign(self._ctx, _ptr(feat), feat.shape[0], _ptr(cent), cent.shape[0], assign_mode,
                                      _ptr(idx), _ptr(err), C.byref(st)), "yupvq_assign")
        return idx, err, st

    def seed_ldr(self, px, ch, k, seed):
        px = _c(px, np.uint8).reshape(-1, 16 * ch)
        out = np.empty(k, np.uint32)
        rounds = C.c_uint32(0)
        self._ck(self._L.yupvq_seed_ldr(self._ctx, _ptr(px), px.shape[0], ch, k, seed, _ptr(out), C.byref(rounds)),
                 "yupvq_seed_ldr")
        return out, int(rounds.value)

    def seed_hdr(self, feat, k, seed):
        feat = _c(feat, np.float32).reshape(-1, 64)
        out = np.empty(k, np.uint32)
        rounds = C.c_uint32(0)
        self._ck(self._L.yupvq_seed_hdr(self._ctx, _ptr(feat), feat.shape[0], k, seed, _ptr(out), C.byref(rounds)),
                 "yupvq_seed_hdr")
        return out, int(rounds.value)

    def update_centroids(self, feat, assign, errors, cent):
        feat = _c(feat, np.float32).reshape(-1, 64)
        assign = _c(assign, np.uint32)
        errors = _c(errors, np.float32)
        cent = np.array(cent, dtype=np.float32, copy=True).reshape(-1, 64)
        counts = np.empty(cent.shape[0], np.uint32)
        self._ck(self._L.yupvq_update_centroids(self._ctx, _ptr(feat), feat.shape[0], _ptr(assign), _ptr(errors),
                                                cent.shape[0], _ptr(cent), _ptr(counts)), "yupvq_update_centroids")
        return cent, counts

    def gather_codebook(self, codebook, indices):
        codebook = _c(codebook, np.uint8)
        n_entries, bs = codebook.shape
        indices = _c(indices, np.uint32)
        out = np.empty((indices.size, bs), np.uint8)
        self._ck(self._L.yupvq_gather_codebook(self._ctx, _ptr(codebook), n_entries, bs, _ptr(indices), indices.size,
                                               _ptr(out)), "yupvq_gather_codebook")
        return out

    def gather_codebook_raw(self, codebook_ptr, n_entries, block_size, indices_ptr, n_blocks, out_ptr):
        """yupvq_gather_codebook on raw (device or host) pointers."""
        self._ck(self._L.yupvq_gather_codebook(self._ctx, _vp(codebook_ptr), n_entries, block_size, _vp(indices_ptr), n_blocks,
                                               _vp(out_ptr)), "yupvq_gather_codebook")

    # ---- N4: whole-texture BCn pre-pass on the GPU
    def bcn_encode_image(self, img, fmt, swap_rb=None, alpha_threshold=128):
        """img: (H, W, ch) uint8 -> (BCn bytes as (n_blocks, blockSize) uint8, kernel ms).  swap_rb defaults to the
        reference's VQ-path behaviour (3-channel sources are read red/blue-swapped by Compressonator)."""
        img = _c(img, np.uint8)
        h, w, ch = img.shape
        swap = (ch == 3) if swap_rb is None else bool(swap_rb)
        bs = 8 if fmt in (BC["BC1"], BC["BC4"]) else 16
        out = np.empty((((w + 3) // 4) * ((h + 3) // 4), bs), np.uint8)
        ms = C.c_float(0)
        self._ck(self._L.yupvq_bcn_encode_image(self._ctx, _ptr(img), w, h, ch, fmt, int(swap), alpha_threshold, _ptr(out), C.byref(ms)),
                 "yupvq_bcn_encode_image")
        return out, ms.value

    def bcn_encode_image_raw(self, img_ptr, w, h, ch, fmt, swap_rb, alpha_threshold, out_ptr):
        """yupvq_bcn_encode_image on raw (device or host) pointers; returns the kernel's CUDA-event time in ms."""
        ms = C.c_float(0)
        self._ck(self._L.yupvq_bcn_encode_image(self._ctx, _vp(img_ptr), w, h, ch, fmt, int(swap_rb), alpha_threshold, _vp(out_ptr),
                                                C.byref(ms)), "yupvq_bcn_encode_image")
        return ms.value

    def compress_image(self, img, fmt, cfg, swap_rb=None, alpha_threshold=128, want_bc=False):
        """Pre-pass + VQ in one call: -> (pixel centroids, indices, stats, BCn encode ms, BCn bytes or None)."""
        img = _c(img, np.uint8)
        h, w, ch = img.shape
        swap = (ch == 3) if swap_rb is None else bool(swap_rb)
        n = ((w + 3) // 4) * ((h + 3) // 4)
        bs = 8 if fmt in (BC["BC1"], BC["BC4"]) else 16
        bc = np.empty((n, bs), np.uint8) if want_bc else None
        cent = np.empty((cfg.codebook_size, 16 * ch), np.uint8)
        idx = np.empty(n, np.uint32)
        st = Stats()
        ms = C.c_float(0)
        self._ck(self._L.yupvq_compress_image(self._ctx, _ptr(img), w, h, ch, fmt, int(swap), alpha_threshold, C.byref(cfg),
                                              _ptr(bc) if want_bc else None, _ptr(cent), _ptr(idx), C.byref(st), C.byref(ms)),
                 "yupvq_compress_image")
        return cent, idx, st, ms.value, bc

    # ---- multi-GPU
    def comm_unique_id(self):
        buf = np.zeros(128, np.uint8)
        self._ck(self._L.yupvq_comm_unique_id(self._ctx, _ptr(buf)), "yupvq_comm_unique_id")
        return buf

    def comm_init(self, uid, n_ranks, rank):
        uid = _c(uid, np.uint8)
        self._ck(self._L.yupvq_comm_init(self._ctx, _ptr(uid), n_ranks, rank), "yupvq_comm_init")
