"""A/B timing of the tensor front-end inside real encode calls (set YUPVQ_LIB to compare library builds;
AB_BATCH=T times the lock-step batch of T textures on one GPU instead of the single-texture call)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yuptexture_b200 import binding, synth
enc = binding.Encoder(0)
T = int(os.environ.get("AB_BATCH", "0"))
size = int(os.environ.get("AB_SIZE", "4096"))
texs = [synth.bc1_encode(synth.ldr_rgb(size, size, seed=1000 + t)) for t in range(max(1, T))]
n = len(texs[0])
cfg = binding.make_config(quality=0.8, max_iterations=32, seed=42)
tag = os.environ.get("YUPVQ_LIB", "current") + (f" batch{T}" if T else "")
for rep in range(4):
    enc.profile(True)
    if T:
        _, _, st = enc.encode_bcn_batch_sharded(texs, n, 0, n, 1, 3, cfg)
    else:
        _, _, st = enc.encode_bcn(texs[0], n, 1, 3, cfg)
    ms, k = enc.profile_read()
    print(tag, "front-end ms/launch %.4f (%d launches) seed %.3f decode %.3f lloyd %.3f quant %.3f total %.3f passes %d" %
          (ms / max(1, k), k, st.ms_seed, st.ms_decode, st.ms_lloyd, st.ms_quantize, st.ms_total, st.lloyd_passes))
enc.close()
