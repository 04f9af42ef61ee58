"""Per-phase timings and re-rank statistics of the config-2 path at several Lloyd pass counts."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yuptexture_b200 import binding, synth
enc = binding.Encoder(0)
img = synth.ldr_rgb(4096, 4096, seed=1000)
bc = synth.bc1_encode(img)
n = len(bc)
prev = None
for iters in [int(x) for x in os.environ.get('LLOYD_ITERS', '1,2,4,8,16,32').split(',')]:
    cfg = binding.make_config(quality=0.8, max_iterations=iters, seed=42)
    for _ in range(2):
        cent, idx, st = enc.encode_bcn(bc, n, 1, 3, cfg)
    d = st.as_dict()
    print(f"iters={iters:2d} passes={d['lloyd_passes']} lloyd={d['ms_lloyd']:.3f} ms ({d['ms_lloyd']/max(1,d['lloyd_passes']):.3f}/pass) "
          f"seed={d['ms_seed']:.3f} decode={d['ms_decode']:.3f} quant={d['ms_quantize']:.3f} total={d['ms_total']:.3f} "
          f"pairs={d['rerank_blocks']} ({d['rerank_blocks']/max(1,d['lloyd_passes'])/n:.3f}/block/pass) overflow={d['rerank_overflow']}", flush=True)
enc.close()
