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


# ---- the exchange step of a sharded pass (csrc/tail.cu, XchgArgs in csrc/kernels.h) ----------------------
ONE_SHOT_LIMIT_BYTES = 4 << 20


def exchange_plan(words, world):
    """How the in-kernel exchange moves a vector of `words` int64: (one_shot, slice_words).  One shot while the world - 1
    copies of the vector stay within a few MB, else two shots over owner slices of ceil(words / world) (csrc/api.cu
    xchg_args)."""
    one_shot = (world - 1) * words * 8 <= ONE_SHOT_LIMIT_BYTES
    return one_shot, (words + world - 1) // world


def exchange_push(partial, world, rank, one_shot, slice_words):
    """What rank `rank` stores into the peers' staging slots: a list of `world` arrays, entry j going to rank j's slot for
    `rank` (one shot: the whole vector to everyone else; two shots: slice j to its owner j)."""
    partial = np.asarray(partial, np.int64)
    if one_shot:
        return [partial if j != rank else partial[:0] for j in range(world)]
    return [partial[j * slice_words:(j + 1) * slice_words] for j in range(world)]


def exchange_reduce_slice(slots):
    """Two shots, phase 1: the owner adds the `world` contributions of its slice (integer adds: order-free)."""
    out = np.zeros_like(slots[0])
    for s in slots:
        out = out + s
    return out


def exchange_sum_on_read(own, slots, rank):
    """One shot: a reader adds its own vector and the other ranks' slots as it goes (csrc/tail.cu ld_acc)."""
    out = np.asarray(own, np.int64).copy()
    for r, s in enumerate(slots):
        if r != rank:
            out = out + s
    return out
