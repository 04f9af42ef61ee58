"""One process, one GPU, but the code path a rank of an N-GPU lock-step batch runs: a communicator of size 1, T textures,
this "rank" owning n / SHARDS blocks of each.  For ncu captures and stage timings of the sharded pass without a
multi-rank launch (the profiler must not wrap one).  The sums cover only the shard, so the codebooks are not the real
ones; the kernels, shapes and launch sequence are.

    python tools/batch_probe.py [T] [SHARDS]
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yuptexture_b200 import binding, synth  # noqa: E402

T = int(sys.argv[1]) if len(sys.argv) > 1 else 2
SH = int(sys.argv[2]) if len(sys.argv) > 2 else 2
enc = binding.Encoder(0)
if os.environ.get("PROBE_COMM", "1") == "1":
    enc.comm_init(enc.comm_unique_id(), 1, 0)
texs = [synth.bc1_encode(synth.ldr_rgb(4096, 4096, seed=1000 + t)) for t in range(T)]
n = len(texs[0])
n_local = n // SH
cfg = binding.make_config(quality=0.8, max_iterations=32, seed=42)
for rep in range(3):
    enc.profile(rep == 2)
    cents, idxs, st = enc.encode_bcn_batch_sharded(texs, n, 0, n_local, 1, 3, cfg)
ms_fe, n_fe = enc.profile_read()
stages, passes = enc.profile_stages()
print("T=%d shard=1/%d rows per launch %d: front-end %.4f ms per launch (%d), stages %s, total %.3f ms" % (
    T, SH, T * n_local, ms_fe / max(1, n_fe), n_fe, {k: round(v, 3) for k, v in stages.items()}, st.ms_total))
enc.close()
