"""GPU search (one-off helper for tools/variant_hash.py): which small synthetic textures make the Lloyd loop hit empty
clusters AFTER pass 0 (so the reseed sees non-zero errors) on the tensor path?  Prints reseeds per candidate."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yuptexture_b200 import binding, synth  # noqa: E402


palette_texture = synth.palette_rgb


def posterised(w, h, seed, step):
    img = synth.ldr_rgb(w, h, seed=seed)
    return ((img // step) * step).astype(np.uint8)


enc = binding.Encoder(0)
w, h = 1024, 512
cands = []
for seed in range(1, 5):
    for m, jit in ((40, 0.02), (80, 0.01), (200, 0.0), (30, 0.2)):
        cands.append((f"palette m={m} jitter={jit} seed={seed}", palette_texture(w, h, m, seed, jit)))
    for step in (32, 64):
        cands.append((f"posterised step={step} seed={seed}", posterised(w, h, seed, step)))
# tiny clusters: few blocks per centroid, where a centroid can lose its last member after pass 0
small = []
for seed in range(1, 4):
    small.append((f"ldr 512x256 seed={seed}", synth.ldr_rgb(512, 256, seed=seed)))
    small.append((f"palette m=600 seed={seed}", palette_texture(512, 256, 600, seed, 0.0)))
    small.append((f"palette m=1100 seed={seed}", palette_texture(512, 256, 1100, seed, 0.0)))
for name, img in small:
    bc = synth.bc1_encode(img)
    for k in (512, 1024):
        for s in (3, 4, 5):
            cfg = binding.make_config(codebook=k, max_iterations=32, seed=s, assign_mode=binding.ASSIGN_TENSOR)
            _, _, st = enc.encode_bcn(bc, len(bc), 1, 3, cfg)
            print(f"{name:40s} n={len(bc)} K={k:4d} seed={s} passes={st.lloyd_passes:2d} reseeds={st.reseeds}", flush=True)
for name, img in cands:
    bc = synth.bc1_encode(img)
    px = enc.decode_bcn(bc, 1, 3)
    distinct = len(np.unique(px, axis=0))
    for k in (64, 128):
        for s in (3, 4):
            cfg = binding.make_config(codebook=k, max_iterations=12, seed=s, assign_mode=binding.ASSIGN_TENSOR)
            _, _, st = enc.encode_bcn(bc, len(bc), 1, 3, cfg)
            print(f"{name:40s} distinct={distinct:6d} K={k:4d} seed={s} passes={st.lloyd_passes:2d} reseeds={st.reseeds}", flush=True)
enc.close()
