"""GPU bring-up check of the tcgen05 assignment: tensor engine vs exact engine vs oracle, plus timing."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yuptexture_b200 import binding, synth
from oracle import orcbind

enc = binding.Encoder(0)
ok = True


def check(name, feat, cent, full_oracle=True):
    global ok
    i_e, e_e, st_e = enc.assign(feat, cent, assign_mode=1)
    i_t, e_t, st_t = enc.assign(feat, cent, assign_mode=2)
    bad = int((i_e != i_t).sum())
    bad_e = int((e_e.view(np.uint32) != e_t.view(np.uint32)).sum())
    msg = f"{name}: n={len(feat)} K={len(cent)} idx_mismatch={bad} err_mismatch={bad_e} exact_ms={st_e.ms_total:.3f} tensor_ms={st_t.ms_total:.3f}"
    if full_oracle:
        i_r, e_r = orcbind.assign(feat, cent)
        msg += f" exact_vs_oracle={int((i_r != i_e).sum())}"
    print(msg, flush=True)
    if bad or bad_e:
        ok = False
        w = np.nonzero(i_e != i_t)[0][:8]
        print("   first mismatches:", w, i_e[w], i_t[w], e_e[w], e_t[w], flush=True)


rng = np.random.default_rng(0)
# structured: real texture + k-means++ seeds as centroids
img = synth.ldr_rgb(1024, 1024, seed=1000)
px = orcbind.decode_bcn(synth.bc1_encode(img), 1, 3)
feat = orcbind.pixels_to_oklab(px, 3)
for k in (512, 256, 1024, 64, 300):
    cpx, _ = orcbind.build_codebook(px[:16384], 3, k, 2, 42)
    cent = orcbind.pixels_to_oklab(cpx, 3)
    check(f"texture k={k}", feat, cent)
check("ragged n", feat[:8192 + 77], cent)
# random pixels: dense near-ties, overflow path
pxr = rng.integers(0, 256, (20000, 48), dtype=np.uint8)
pcr = rng.integers(0, 256, (512, 48), dtype=np.uint8)
check("random", orcbind.pixels_to_oklab(pxr, 3), orcbind.pixels_to_oklab(pcr, 3))
# exact ties
c8 = orcbind.pixels_to_oklab(rng.integers(0, 256, (8, 48), dtype=np.uint8), 3)
centt = np.concatenate([c8] * 16)
featt = np.concatenate([c8] * 1100)
check("ties", featt, centt)
# flat
check("flat", np.tile(c8[:1], (9000, 1)), np.tile(c8[:1], (128, 1)))
# big: config 2 shape
img = synth.ldr_rgb(4096, 4096, seed=1000)
bc = synth.bc1_encode(img)
cfg = binding.make_config(quality=0.8, max_iterations=4, seed=42)
cent_px, idx, st = enc.encode_bcn(bc, len(bc), 1, 3, cfg)
print("encode(auto) stats:", st.as_dict(), flush=True)
cfg.assign_mode = 1
cent_px_e, idx_e, st_e = enc.encode_bcn(bc, len(bc), 1, 3, cfg)
print("encode(exact) stats:", st_e.as_dict(), flush=True)
print("centroids equal:", np.array_equal(cent_px, cent_px_e), "idx mismatches:", int((idx != idx_e).sum()), flush=True)
if not np.array_equal(cent_px, cent_px_e) or (idx != idx_e).any():
    ok = False
pxb = enc.decode_bcn(bc, 1, 3)
featb = enc.pixels_to_oklab(pxb, 3)
centb = enc.pixels_to_oklab(cent_px, 3)
check("config2", featb, centb, full_oracle=False)
print("TC_CHECK", "PASS" if ok else "FAIL", flush=True)
enc.close()
sys.exit(0 if ok else 1)
