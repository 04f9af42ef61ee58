"""A/B timing of the tensor front-end inside real encode calls (set YUPVQ_LIB to compare library builds)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yuptexture_b200 import binding, synth
enc = binding.Encoder(0)
bc = synth.bc1_encode(synth.ldr_rgb(4096, 4096, seed=1000))
n = len(bc)
cfg = binding.make_config(quality=0.8, max_iterations=32, seed=42)
for rep in range(4):
    enc.profile(True)
    _, _, st = enc.encode_bcn(bc, n, 1, 3, cfg)
    ms, k = enc.profile_read()
    print(os.environ.get("YUPVQ_LIB", "current"), "front-end ms/launch %.4f (%d launches) lloyd %.3f total %.3f" % (ms / max(1, k), k, st.ms_lloyd, st.ms_total))
enc.close()
