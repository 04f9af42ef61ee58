"""Prints one hash over the outputs of a fixed set of encodes (single texture BC1 / BC3, lock-step batch).  The implementation
switches of csrc/host_util.h (environment, read once per process) must not change it: tests/test_gpu_pipeline.py runs this
script once per switch and compares the lines."""
import hashlib
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yuptexture_b200 import binding, synth  # noqa: E402

enc = binding.Encoder(0)
h = hashlib.sha256()
rgb = synth.ldr_rgb(1024, 1024, seed=31)
bc1 = synth.bc1_encode(rgb)
cfg = binding.make_config(codebook=256, max_iterations=12, seed=5)
cent, idx, st = enc.encode_bcn(bc1, len(bc1), 1, 3, cfg)
h.update(cent.tobytes()); h.update(idx.tobytes()); h.update(bytes([st.lloyd_passes]))
bc3 = synth.bc3_encode(synth.ldr_rgba(512, 512, seed=32))
cent, idx, st = enc.encode_bcn(bc3, len(bc3), 3, 4, binding.make_config(codebook=128, max_iterations=8, seed=6))
h.update(cent.tobytes()); h.update(idx.tobytes())
texs = [synth.bc1_encode(synth.ldr_rgb(512, 512, seed=40 + t)) for t in range(3)]
cents, idxs, st = enc.encode_bcn_batch_sharded(texs, len(texs[0]), 0, len(texs[0]), 1, 3, binding.make_config(codebook=128, max_iterations=8, seed=7))
for c, i in zip(cents, idxs):
    h.update(np.ascontiguousarray(c).tobytes()); h.update(np.ascontiguousarray(i).tobytes())
# duplicate-heavy texture on the tensor path: two distinct blocks, K = 64 -> 62 empty clusters on every pass, so
# the fused tail's rare path runs (errors of all rows, grid-wide first-maximum arg-max across the CTA stripes -- every error
# ties at 0, so the FIRST stripe's first row must win -- reseed, operand rows again) and must equal the separate
# exact_errors / reseed_empty kernels that YUPVQ_NO_TAIL_FUSE selects
chk = synth.bc1_encode(synth.checker_rgb(1024, 512))
cent, idx, st = enc.encode_bcn(chk, len(chk), 1, 3, binding.make_config(codebook=64, max_iterations=6, seed=8, assign_mode=binding.ASSIGN_TENSOR))
assert st.reseeds > 0, "the checker case no longer produces empty clusters"
h.update(cent.tobytes()); h.update(idx.tobytes()); h.update(bytes([st.lloyd_passes]))
# ... and the same path with NON-zero errors: 1100 distinct blocks, 8 blocks per centroid at K = 1024 -- clusters run empty
# after the first pass, so the arg-max works on real distances (a different winner per empty cluster)
pal = synth.bc1_encode(synth.palette_rgb(512, 256, 1100, 1))
cent, idx, st = enc.encode_bcn(pal, len(pal), 1, 3, binding.make_config(codebook=1024, max_iterations=8, seed=3, assign_mode=binding.ASSIGN_TENSOR))
assert st.reseeds > 0 and st.lloyd_passes > 2, "the palette case no longer loses clusters after pass 0"
h.update(cent.tobytes()); h.update(idx.tobytes()); h.update(bytes([st.lloyd_passes]))
print("VARIANT_HASH", h.hexdigest())
enc.close()
