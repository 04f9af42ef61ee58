"""Times the tensor assignment on the config-2 shape (for ncu launch lists / debug variants)."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yuptexture_b200 import binding
enc = binding.Encoder(0)
rng = np.random.default_rng(0)
n, k = 1 << 20, int(os.environ.get("TC_K", "512"))
base = rng.normal(0, 0.1, (k, 64)).astype(np.float32) + 0.5
feat = (base[rng.integers(0, k, n)] + rng.normal(0, 0.02, (n, 64))).astype(np.float32)
for _ in range(3):
    idx, err, st = enc.assign(feat, base, assign_mode=2)
print("ms_total", st.ms_total, "pairs", st.rerank_blocks)
enc.close()
