"""Times the LDR K-Means++ seeding (config-2 shape) and checks it against the per-round kernels."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yuptexture_b200 import binding, synth
enc = binding.Encoder(0)
img = synth.ldr_rgb(4096, 4096, seed=1000)
bc = synth.bc1_encode(img)
n = len(bc)
cfg = binding.make_config(quality=0.8, max_iterations=0, seed=42)
for _ in range(3):
    cent, idx, st = enc.encode_bcn(bc, n, 1, 3, cfg)
print("seed ms", st.ms_seed, "rounds", st.seed_rounds, "launches", st.gpu_launches)
np.save(os.environ.get("SEED_OUT", "/tmp/seed_cent.npy"), cent)
enc.close()
