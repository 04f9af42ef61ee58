"""Phase times of the other BASELINE configs on one GPU (config 3: 4096^2 RGBA as BC3; config 4: 4096x2048 HDR, K = 1024)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yuptexture_b200 import binding, synth
enc = binding.Encoder(0)

def show(tag, st):
    print(tag, "decode %.3f seed %.3f lloyd %.3f (%d passes, %.3f ms/pass) finalize %.3f quant %.3f total %.3f ms; rerank pairs %d overflow %d" %
          (st.ms_decode, st.ms_seed, st.ms_lloyd, st.lloyd_passes, st.ms_lloyd / max(1, st.lloyd_passes), st.ms_finalize, st.ms_quantize,
           st.ms_total, st.rerank_blocks, st.rerank_overflow))

bc3 = synth.bc3_encode(synth.ldr_rgba(4096, 4096, seed=5))
cfg = binding.make_config(quality=0.8, max_iterations=32, seed=43)
for _ in range(3):
    _, _, st = enc.encode_bcn(bc3, len(bc3), 3, 4, cfg)
show("config3 BC3 RGBA 4096^2 K=512:", st)
img = synth.hdr_rgb(4096, 2048, seed=11)
cfg = binding.make_config(quality=1.0, max_iterations=64, seed=44)
for _ in range(3):
    _, _, st = enc.encode_hdr(img, 4096, 2048, 3, cfg)
show("config4 HDR 4096x2048 K=1024:", st)
cfg = binding.make_config(quality=0.8, max_iterations=32, seed=44)
for _ in range(2):
    _, _, st = enc.encode_hdr(img, 4096, 2048, 3, cfg)
show("        HDR 4096x2048 K=512:", st)
enc.close()
