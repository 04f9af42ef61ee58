"""Times the SAD-metric path (vq_Metric == SAD) on the config-2 texture: seeding + Lloyd on pixel bytes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yuptexture_b200 import binding, synth
enc = binding.Encoder(0)
bc = synth.bc1_encode(synth.ldr_rgb(4096, 4096, seed=1000))
n = len(bc)
cfg = binding.make_config(quality=0.8, max_iterations=32, metric=binding.METRIC_SAD, seed=42)
for _ in range(3):
    cent, idx, st = enc.encode_bcn(bc, n, 1, 3, cfg)
print("SAD path: seed %.3f lloyd %.3f ms, %d passes -> %.3f ms/pass; quantize %.3f" % (
    st.ms_seed, st.ms_lloyd, st.lloyd_passes, st.ms_lloyd / max(1, st.lloyd_passes), st.ms_quantize))
enc.close()
