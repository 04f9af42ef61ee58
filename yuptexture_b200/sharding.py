"""Host-side partitioning of the VQ path over one process per GPU.

Blocks shard contiguously: rank r owns [first, first + n_local) of every texture; the only data-path
exchange is the per-pass all-reduce of the fixed-point [K x 64 sums | K counts | changed] words, which
is associative (int64), so any rank count gives the same codebook.  Seeding is texture-parallel in a
lock-step batch: texture t is seeded by rank t % world.
"""
import numpy as np


def shard_range(n_blocks, world, rank):
    """Contiguous block range of `rank`: (first, n_local).  The ranges tile [0, n_blocks)."""
    per = (n_blocks + world - 1) // world
    first = min(rank * per, n_blocks)
    return first, min(per, n_blocks - first)


def seeding_owner(texture, world):
    """Rank that decodes texture `texture` whole and runs its K-Means++ seeding (phase A)."""
    return texture % world


def fixed_shift(n_total, max_abs):
    """Fixed-point scale of the centroid sums (csrc/api.cu fixed_shift): n_total values of magnitude
    <= max_abs summed as round(v * 2^shift) must stay below 2^62."""
    bits_n = 1
    while (1 << bits_n) < n_total and bits_n < 40:
        bits_n += 1
    e = 0
    if np.isfinite(max_abs) and max_abs > 0:
        e = int(np.frexp(np.float32(max_abs))[1])
    elif not np.isfinite(max_abs):
        e = 128
    return max(-60, min(62 - bits_n - max(e, 0), 40))


def partial_sums(features, assign, k, shift):
    """One rank's [K*64 sums | K counts] int64 words for its shard (what launch_accumulate produces)."""
    feats = np.asarray(features, np.float32).reshape(-1, 64)
    fx = np.rint(feats.astype(np.float64) * (2.0 ** shift)).astype(np.int64)
    acc = np.zeros(k * 65, np.int64)
    np.add.at(acc[:k * 64].reshape(k, 64), assign, fx)
    np.add.at(acc[k * 64:], assign, 1)
    return acc
