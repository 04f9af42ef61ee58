"""CPU simulation of exact k-means bounds (Hamerly row skipping, Yinyang group filtering) on the bench texture.

Answers one question before any kernel is written: what fraction of rows / (row, centroid-group) pairs would an exact
bounds scheme still have to scan in each Lloyd pass of BASELINE config 2 (4096^2 BC1, K = 512)?  Features and K-Means++
seeds come from the oracle (test infrastructure; this tool is not product code), the Lloyd loop is float32 torch on
the CPU.  Prints one line per pass:  moved rows | rows Hamerly must scan | rows / pairs a G-group Yinyang filter must scan.

    python tools/bounds_sim.py 4096 8
"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import orcbind  # noqa: E402
from yuptexture_b200 import synth  # noqa: E402


def main():
    S = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
    G = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    K = 512
    t0 = time.time()
    bc = synth.bc1_encode(synth.ldr_rgb(S, S, seed=1000))
    px = orcbind.decode_bcn(bc, 1, 3)
    feat = torch.from_numpy(orcbind.pixels_to_oklab(px, 3).reshape(len(px), 64).astype(np.float32))
    seeds, _ = orcbind.seed_ldr(px, 3, K, 42)
    X = feat * torch.tensor([2., 1., 1., 1.]).repeat(16)
    N = X.shape[0]
    C = X[torch.from_numpy(seeds.astype(np.int64))].clone()
    xn = (X * X).sum(1)
    grp = torch.arange(K) // (K // G)
    print("N=%d K=%d G=%d (setup %.1f s)" % (N, K, G, time.time() - t0), flush=True)
    CH = 1 << 17

    def scan(Cm, a_prev):
        """full distances chunk by chunk -> nearest, its distance, second nearest, per-group minima excluding the nearest,
        and the distance to the previous centroid"""
        cn = (Cm * Cm).sum(1)
        a = torch.empty(N, dtype=torch.long)
        d1 = torch.empty(N)
        d2 = torch.empty(N)
        dg = torch.empty(N, G)
        dp = torch.empty(N)
        for s in range(0, N, CH):
            e = min(N, s + CH)
            D = (xn[s:e, None] - 2 * X[s:e] @ Cm.T + cn[None, :]).clamp_(min=0).sqrt_()
            dp[s:e] = D[torch.arange(e - s), a_prev[s:e]]
            d1[s:e], a[s:e] = D.min(1)
            D[torch.arange(e - s), a[s:e]] = float("inf")
            d2[s:e] = D.min(1).values
            dg[s:e] = D.view(e - s, G, K // G).min(2).values
        return a, d1, d2, dg, dp

    a_prev = torch.zeros(N, dtype=torch.long)
    tot_h = tot_r = tot_p = 0.0
    for it in range(32):
        a, d1, d2, dg, dp = scan(C, a_prev)
        if it > 0:
            drift = (C - C_old).norm(dim=1)
            gd = torch.stack([drift[grp == g].max() for g in range(G)])
            lg2 = lg - gd[None, :]
            lh2 = lh - drift.max()
            need_h = dp >= lh2                    # after tightening the upper bound to the exact own-centroid distance
            need_g = dp[:, None] >= lg2
            rows_g = need_g.any(1)
            assert (need_h | (a == a_prev)).all() and (rows_g | (a == a_prev)).all()
            fh, fr, fp = need_h.float().mean().item(), rows_g.float().mean().item(), need_g.float().mean().item()
            tot_h += fh; tot_r += fr; tot_p += fp
            print("pass %2d moved %.4f | Hamerly rows %.3f | G=%d rows %.3f pairs %.3f" % (
                it, (a != a_prev).float().mean().item(), fh, G, fr, fp), flush=True)
            lh = torch.where(need_h, d2, lh2)
            lg = torch.where(rows_g[:, None], dg, lg2)
        else:
            lh, lg = d2, dg
        C_old = C.clone()
        cnt = torch.bincount(a, minlength=K).float()
        sums = torch.zeros(K, 64).index_add_(0, a, X)
        nz = cnt > 0
        C[nz] = sums[nz] / cnt[nz][:, None]
        a_prev = a
    print("mean over passes 1..31: Hamerly rows %.3f | Yinyang rows %.3f pairs %.3f" % (tot_h / 31, tot_r / 31, tot_p / 31))


if __name__ == "__main__":
    main()
