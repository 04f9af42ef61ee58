// K2 / K2h: K-Means++ D^2 seeding, one launch per round, no host round trip.
// Reference: src/vq_encoder.cpp:287-313 (LDR: SAD^2 on uint8 pixel blocks) and :526-554 (HDR:
// weighted L2 on Oklab features).  Each round fuses
//   minDist[j] = min(minDist[j], dist(block_j, seed_{i-1}))      (the reference's omp loop)
//   S = sum_j minDist[j]                                         (warp-shuffle + CTA reduction)
//   p = u_i * S;  first j with prefix(j) >= p                    (two-level prefix search, done by
//                                                                 the last CTA to finish)
// LDR values are integer-valued floats (< 2^29) and N * 2^29 < 2^53, so the uint64 sums/prefixes
// here equal the reference's double sums exactly and the chosen seeds are bit-identical for the
// same mt19937 draws.  HDR sums are doubles reduced in a fixed tree.
// Roofline: HBM/L2-bound, (16*ch + 8) B per block per round; additionally latency-bound by K-1
// dependent launches.
#include <cfloat>
#include <cstdio>
#include <cstdlib>
#include <type_traits>

#include "common.cuh"
#include "host_util.h"
#include "kernels.h"

namespace yupvq {

namespace {

constexpr int kThreads = 256;
constexpr int kMaxGrid = YUPVQ_NUM_SMS * 8;

// PRMT with raw selector nibbles (all in 0..3 here, upper 16 bits ignored by the hardware): __byte_perm would mask the
// selector first.  VABSDIFF4 with the accumulate operand spelled out: one instruction per four byte pairs.
__device__ __forceinline__ uint32_t prmt_lo(uint32_t a, uint32_t sel) {
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(0u), "r"(sel));
    return d;
}
__device__ __forceinline__ uint32_t sad4_acc(uint32_t a, uint32_t b, uint32_t acc) {
    uint32_t d;
    asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(acc));
    return d;
}


__global__ void seed_init_kernel(float* min_dist, size_t n, unsigned int* counter, uint32_t* status, uint32_t* seed_idx,
                                 uint32_t first_idx) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) min_dist[i] = FLT_MAX;
    if (i == 0) {
        *counter = 0;
        status[0] = 0;
        status[1] = 0;
        seed_idx[0] = first_idx;
    }
}

template <typename T>
__device__ __forceinline__ T block_sum(T v, T* smem) {
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) smem[wid] = v;
    __syncthreads();
    T r = (threadIdx.x < kThreads / 32) ? smem[threadIdx.x] : (T)0;
    if (wid == 0) {
        r = warp_sum(r);
        if (lane == 0) smem[0] = r;
    }
    __syncthreads();
    r = smem[0];
    __syncthreads();
    return r;
}

// exclusive prefix over the CTA's per-thread values (kThreads threads); returns the exclusive prefix
// of this thread; *total receives the CTA sum.
template <typename T>
__device__ __forceinline__ T block_exclusive_scan(T v, T* smem, T* total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    T incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const T t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) smem[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        T w = (lane < kThreads / 32) ? smem[lane] : (T)0;
        T wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const T t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += t;
        }
        if (lane < kThreads / 32) smem[lane] = wi - w;  // exclusive prefix of warp sums
        if (lane == kThreads / 32 - 1) smem[kThreads / 32] = wi;
    }
    __syncthreads();
    const T r = smem[wid] + (incl - v);
    *total = smem[kThreads / 32];
    __syncthreads();
    return r;
}

// The weighted-sample search run by the last CTA.  ACC is uint64 (LDR, exact) or double (HDR).
// value(j) returns minDist[j] converted to ACC.
template <typename ACC>
__device__ void pick_sample(const SeedState& st, size_t n, size_t chunk, uint32_t round, ACC* smem_acc,
                            unsigned int* smem_u32) {
    const ACC* partial = reinterpret_cast<const ACC*>(st.partial);
    const int grid = st.grid;
    // ---- level 1: which CTA chunk holds the sample
    const int per = (grid + kThreads - 1) / kThreads;
    const int p0 = threadIdx.x * per;
    ACC local = (ACC)0;
    for (int k = 0; k < per; ++k)
        if (p0 + k < grid) local += __ldcg(&partial[p0 + k]);
    ACC total;
    ACC excl = block_exclusive_scan<ACC>(local, smem_acc, &total);
    if (!((double)total > 0.0)) {  // reference: if (current_sum <= 0) fill the rest with seed 0 and stop
        if (threadIdx.x == 0) {
            st.status[0] = round;
            st.seed_idx[round] = st.seed_idx[0];
        }
        return;
    }
    const double p = __dadd_rn(__dmul_rn(st.canon[round - 1], (double)total), 0.0);
    if (threadIdx.x == 0) *smem_u32 = 0xffffffffu;
    __syncthreads();
    {
        ACC cum = excl;
        for (int k = 0; k < per; ++k)
            if (p0 + k < grid) {
                cum += __ldcg(&partial[p0 + k]);
                if ((double)cum >= p) {
                    atomicMin(smem_u32, (unsigned int)(p0 + k));
                    break;
                }
            }
    }
    __syncthreads();
    unsigned int cta = *smem_u32;
    __syncthreads();
    if (cta == 0xffffffffu) cta = (unsigned int)(grid - 1);  // unreachable: cum(last) == total >= p
    // exclusive prefix of the chosen chunk, recomputed in index order (same association for every thread)
    ACC base = (ACC)0;
    {
        ACC part = (ACC)0;
        for (unsigned int k = threadIdx.x; k < cta; k += kThreads) part += __ldcg(&partial[k]);
        // for doubles a strided sum is not the sequential one; the difference is ~1e-16 relative and
        // only matters if it moves p across a bucket boundary.  Exact for uint64.
        base = block_sum<ACC>(part, smem_acc);
    }
    // ---- level 2: which block inside the chunk
    const size_t begin = (size_t)cta * chunk;
    const size_t end = begin + chunk < n ? begin + chunk : n;
    const size_t len = end - begin;
    const size_t seg = (len + kThreads - 1) / kThreads;
    const size_t s0 = begin + (size_t)threadIdx.x * seg;
    ACC segsum = (ACC)0;
    for (size_t j = s0; j < s0 + seg && j < end; ++j) segsum += (ACC)__ldcg(&st.min_dist[j]);
    ACC ctot;
    ACC sexcl = block_exclusive_scan<ACC>(segsum, smem_acc, &ctot);
    if (threadIdx.x == 0) *smem_u32 = 0xffffffffu;
    __syncthreads();
    {
        ACC cum = base + sexcl;
        for (size_t j = s0; j < s0 + seg && j < end; ++j) {
            cum += (ACC)__ldcg(&st.min_dist[j]);
            if ((double)cum >= p) {
                atomicMin(smem_u32, (unsigned int)(j - begin));
                break;
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int off = *smem_u32;
        if (off == 0xffffffffu) off = (unsigned int)(len - 1);
        st.seed_idx[round] = (uint32_t)(begin + off);
        st.status[1] = st.status[1] + 1;
    }
}

template <int CH>
__global__ void __launch_bounds__(kThreads) seed_round_ldr_kernel(SeedState st, const uint4* __restrict__ planes,
                                                                  size_t stride, size_t n, size_t chunk,
                                                                  uint32_t round) {
    __shared__ uint4 seed_px[4];
    __shared__ unsigned long long red[kThreads / 32 + 1];
    __shared__ unsigned int s_u32;
    if (st.status[0] != 0) {  // degenerate texture: every remaining seed is seed 0 (reference :299-302)
        if (blockIdx.x == 0 && threadIdx.x == 0) st.seed_idx[round] = st.seed_idx[0];
        return;
    }
    const uint32_t prev = st.seed_idx[round - 1];
    if (threadIdx.x < CH) seed_px[threadIdx.x] = planes[threadIdx.x * stride + prev];
    __syncthreads();
    uint4 sp[CH];
#pragma unroll
    for (int q = 0; q < CH; ++q) sp[q] = seed_px[q];

    const size_t begin = (size_t)blockIdx.x * chunk;
    const size_t end = begin + chunk < n ? begin + chunk : n;
    unsigned long long local = 0;
    for (size_t j = begin + threadIdx.x; j < end; j += kThreads) {
        unsigned int sad = 0;
#pragma unroll
        for (int q = 0; q < CH; ++q) {
            const uint4 v = planes[q * stride + j];
            sad = sad4_acc(v.x, sp[q].x, sad);
            sad = sad4_acc(v.y, sp[q].y, sad);
            sad = sad4_acc(v.z, sp[q].z, sad);
            sad = sad4_acc(v.w, sp[q].w, sad);
        }
        const float d = (float)sad;                       // BlockDistanceSAD returns float(sad)
        const float m = fminf(__fmul_rn(d, d), st.min_dist[j]);  // std::min(d * d, minDistSq[j])
        st.min_dist[j] = m;
        local += (unsigned long long)m;                   // integer-valued, < 2^29: exact
    }
    const unsigned long long csum = block_sum<unsigned long long>(local, red);
    __shared__ bool is_last;
    if (threadIdx.x == 0) {
        st.partial[blockIdx.x] = csum;
        __threadfence();
        const unsigned int ticket = atomicAdd(st.counter, 1u);
        is_last = (ticket == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    pick_sample<unsigned long long>(st, n, chunk, round, red, &s_u32);
    if (threadIdx.x == 0) *st.counter = 0;
}

__global__ void __launch_bounds__(kThreads) seed_round_hdr_kernel(SeedState st, const float* __restrict__ feat,
                                                                  size_t n, size_t chunk, uint32_t round) {
    __shared__ float4 seed_f[16];
    __shared__ double red[kThreads / 32 + 1];
    __shared__ unsigned int s_u32;
    if (st.status[0] != 0) {
        if (blockIdx.x == 0 && threadIdx.x == 0) st.seed_idx[round] = st.seed_idx[0];
        return;
    }
    const uint32_t prev = st.seed_idx[round - 1];
    if (threadIdx.x < 16) seed_f[threadIdx.x] = reinterpret_cast<const float4*>(feat)[(size_t)prev * 16 + threadIdx.x];
    __syncthreads();
    const size_t begin = (size_t)blockIdx.x * chunk;
    const size_t end = begin + chunk < n ? begin + chunk : n;
    double local = 0.0;
    for (size_t j = begin + threadIdx.x; j < end; j += kThreads) {
        const float4* x = reinterpret_cast<const float4*>(feat) + j * 16;
        DistAcc acc;
        acc.reset();
#pragma unroll
        for (int i = 0; i < 8; ++i) acc.step(x[2 * i], x[2 * i + 1], seed_f[2 * i], seed_f[2 * i + 1]);
        const float m = fminf(acc.finish(), st.min_dist[j]);  // std::min(d, minDistSq[j])  (d, not d*d)
        st.min_dist[j] = m;
        local += (double)m;
    }
    const double csum = block_sum<double>(local, red);
    __shared__ bool is_last;
    if (threadIdx.x == 0) {
        reinterpret_cast<double*>(st.partial)[blockIdx.x] = csum;
        __threadfence();
        const unsigned int ticket = atomicAdd(st.counter, 1u);
        is_last = (ticket == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    pick_sample<double>(st, n, chunk, round, red, &s_u32);
    if (threadIdx.x == 0) *st.counter = 0;
}

// ================================================================================================
// Persistent variant: ALL K-1 rounds in one cooperative launch (one CTA of 1024 threads per SM).
// A round is latency-, not bandwidth-bound (its working set lives in L2 and partly in shared memory),
// so what matters is the dependency chain  min-update -> global sum -> weighted pick -> next seed:
//   (1) every WARP publishes its partial sum as one tagged 64-bit word [round:16 | sum:48] with a
//       release store (HDR: a {double, tag} pair) -- no CTA barrier on the way;
//   (2) the leader CTA polls all W warp words (five per thread, requested together), scans them, draws
//       p = u * S, finds the owning warp, reads that warp's <= bpw minimum distances from L2 (acquire /
//       release makes them visible) and finds the block;
//   (3) the pick is published as a tagged word that one thread per CTA polls.
// Three L2 round trips per round and no all-to-all polling (148 x 148 pollers on ten cache lines cost
// 8 us per round when tried).  Block ownership is hierarchical and contiguous (warp > 32-block segment
// > lane), so prefix order == block index order, which "first j with prefix >= p" (reference :306-312)
// needs.  LDR sums are uint64 (exact, F12): the search runs on integers, first j with prefix_j >= ceil(p).
// The first `cseg` segments of every warp keep their pixel planes in shared memory for the whole kernel
// (up to ~200 KB per SM), which takes that share of the per-round traffic off the L2.
// ================================================================================================
constexpr int kPThreads = 1024;
constexpr int kPWarps = kPThreads / 32;
constexpr int kPMaxGrid = 160;                    // >= SM count
constexpr int kPPer = kPMaxGrid * kPWarps / kPThreads;   // warp words per leader thread (5)
constexpr unsigned long long kTagShift = 48;
constexpr unsigned long long kPayloadMask = (1ull << kTagShift) - 1ull;
constexpr unsigned long long kStopBit = 1ull << 47;

__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

template <typename T>
__device__ __forceinline__ T warp_incl_scan(T v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const T t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

#ifdef YUPVQ_SEED_TIMING
#define TICK(acc) do { t1 = clock64(); acc += t1 - t0; t0 = t1; } while (0)
#else
#define TICK(acc) do { } while (0)
#endif

// HDR == false: CH uint4 planes per block, SAD^2, uint64 sums.  HDR == true: 64-float rows, weighted L2, double sums.
// Exchange words in st.partial: [0, 2W) warp words (LDR uses the first W), then the pick word.
// TRI (LDR, MDREG): triangle-inequality skip.  SAD is a metric, so with s_j the seed nearest to block j and
// d_j = SAD(j, s_j):  SAD(new, s_j) >= 2 d_j  =>  SAD(j, new) >= d_j  =>  min(d^2, minDistSq[j]) is unchanged and the
// block's pixels need not be read at all.  Every CTA keeps the pixel blocks of all seeds chosen so far in shared
// memory and computes the K' <= 1023 distances new-seed -> earlier-seed at the top of a round (one per thread); a
// lane keeps (min SAD, nearest seed ordinal) packed in one register per block instead of the squared float, which
// is recomputed on demand (float(d) * float(d) is monotone in d, so the minimum of the rounded squares is the rounded
// square of the minimum).  Values, sums and picks are bit-identical to the plain update.
template <bool HDR, int CH, bool MDREG, bool TRI = false>
__global__ void __launch_bounds__(kPThreads, 1)
seed_persistent_kernel(SeedState st, const uint4* __restrict__ planes, size_t stride, const float* __restrict__ feat, size_t n,
                       size_t bpw, int cseg, uint32_t K) {
    using ACC = typename std::conditional<HDR, double, unsigned long long>::type;
    extern __shared__ uint4 s_cache[];            // LDR: [warp][cseg][CH][32] pixel planes
    __shared__ ACC s_scan[kPWarps + 1];
    __shared__ ACC s_target;
    __shared__ unsigned int s_hit;
    __shared__ unsigned long long s_pick;
    __shared__ float4 s_seedf[16];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int W = gridDim.x * kPWarps;
    const size_t gw = (size_t)blockIdx.x * kPWarps + wid;
    const size_t wbase = gw * bpw;
    const size_t wend = wbase + bpw < n ? wbase + bpw : n;  // may be <= wbase: empty warp
    const int nseg = (int)(bpw / 32);
    unsigned long long* wwords = st.partial;
    unsigned long long* pick = st.partial + 2 * kPMaxGrid * kPWarps;
    const bool leader = blockIdx.x == 0;

    if (!HDR) {  // resident pixel planes
        for (int i = 0; i < cseg; ++i) {
            const size_t j = wbase + (size_t)i * 32 + lane;
#pragma unroll
            for (int q = 0; q < CH; ++q)
                s_cache[((wid * cseg + i) * CH + q) * 32 + lane] = j < wend ? ldg_nc_v4(planes + q * stride + j) : make_uint4(0, 0, 0, 0);
        }
        __syncwarp();
    }

    float md[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) md[i] = FLT_MAX;  // seed_init_kernel wrote the same value to min_dist
    uint32_t msnr[8];                             // TRI: [nearest seed ordinal : 16 | min SAD : 16], 0xffff = none yet
#pragma unroll
    for (int i = 0; i < 8; ++i) msnr[i] = 0xffffu;
    uint4* const seeds_s = s_cache + (size_t)cseg * kPWarps * CH * 32;          // TRI: [K][CH] pixel blocks of the seeds
    uint32_t* const tri_s = reinterpret_cast<uint32_t*>(seeds_s + (size_t)K * CH);  // TRI: SAD(new seed, seed q)
    if (TRI && HDR) {   // "seed q is nearest to one of this CTA's blocks" bits (after the distance and index tables)
        uint32_t* used0 = reinterpret_cast<uint32_t*>(s_cache) + 2 * (size_t)K;
        for (uint32_t i = tid; i < (K + 31) / 32; i += kPThreads) used0[i] = 0u;
        __syncthreads();
    }
    // per-lane constants of the update loop
    unsigned int vmask = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i)
        if (i < nseg && wbase + (size_t)i * 32 + lane < wend) vmask |= 1u << i;
    float* const mdp = st.min_dist + wbase + lane;
    const uint4* gp[HDR ? 1 : CH];
#pragma unroll
    for (int q = 0; q < (HDR ? 1 : CH); ++q) gp[q] = planes + q * stride + wbase + lane;
    const uint4* const sc = s_cache + (size_t)wid * cseg * CH * 32 + lane;
    uint32_t seed = st.seed_idx[0];
#ifdef YUPVQ_SEED_TIMING
    long long t_upd = 0, t_poll = 0, t_l2 = 0, t_pick = 0, t_p1 = 0, t_p2 = 0, t0 = 0, t1;
#endif
    for (uint32_t round = 1; round < K; ++round) {
#ifdef YUPVQ_SEED_TIMING
        t0 = clock64();
#endif
        const double canon_u = (leader && tid == 31) ? st.canon[round - 1] : 0.0;  // fetched early: off the critical path
        // ---- the new seed's data
        uint4 sp[HDR ? 1 : CH];
        if (HDR) {
            if (tid < 16) s_seedf[tid] = reinterpret_cast<const float4*>(feat)[(size_t)seed * 16 + tid];
            __syncthreads();
        } else {
#pragma unroll
            for (int q = 0; q < CH; ++q) sp[q] = __ldg(planes + q * stride + seed);
        }
        // ---- min-distance update over this lane's blocks
        ACC local = (ACC)0;
        auto block_dist = [&](int i, size_t j) -> float {
            if (HDR) {
                const float4* x = reinterpret_cast<const float4*>(feat) + j * 16;
                DistAcc acc;
                acc.reset();
#pragma unroll
                for (int k = 0; k < 8; ++k) acc.step(x[2 * k], x[2 * k + 1], s_seedf[2 * k], s_seedf[2 * k + 1]);
                return acc.finish();                                           // HDR seeds on d, not d*d (:536-537)
            } else {
                unsigned int sad = 0;
#pragma unroll
                for (int q = 0; q < CH; ++q) {
                    const uint4 v = i < cseg ? s_cache[((wid * cseg + i) * CH + q) * 32 + lane] : ldg_nc_v4(planes + q * stride + j);
                    sad = sad4_acc(v.x, sp[q].x, sad);
                    sad = sad4_acc(v.y, sp[q].y, sad);
                    sad = sad4_acc(v.z, sp[q].z, sad);
                    sad = sad4_acc(v.w, sp[q].w, sad);
                }
                const float d = (float)sad;                                    // BlockDistanceSAD returns float(sad)
                return __fmul_rn(d, d);
            }
        };
        if (TRI && MDREG && !HDR) {
            const uint32_t o = round - 1;             // ordinal of the seed applied in this round
            if (tid < (int)o) {
                unsigned int sad = 0;
#pragma unroll
                for (int q = 0; q < CH; ++q) {
                    const uint4 v = seeds_s[(size_t)tid * CH + q];
                    sad = sad4_acc(v.x, sp[q].x, sad);
                    sad = sad4_acc(v.y, sp[q].y, sad);
                    sad = sad4_acc(v.z, sp[q].z, sad);
                    sad = sad4_acc(v.w, sp[q].w, sad);
                }
                tri_s[tid] = sad;
            }
            if (tid == 0) {                                          // read from the next round on (barriers in between)
#pragma unroll
                for (int q = 0; q < CH; ++q) seeds_s[(size_t)o * CH + q] = sp[q];
            }
            __syncthreads();
            unsigned int loc32 = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if ((vmask >> i) & 1u) {
                    uint32_t ms = msnr[i] & 0xffffu;
                    const bool skip = ms != 0xffffu && tri_s[msnr[i] >> 16] >= 2u * ms;
                    if (!skip) {
                        unsigned int sad = 0;
                        if (i < cseg) {
#pragma unroll
                            for (int q = 0; q < CH; ++q) {
                                const uint4 v = sc[(i * CH + q) * 32];
                                sad = sad4_acc(v.x, sp[q].x, sad);
                                sad = sad4_acc(v.y, sp[q].y, sad);
                                sad = sad4_acc(v.z, sp[q].z, sad);
                                sad = sad4_acc(v.w, sp[q].w, sad);
                            }
                        } else {
#pragma unroll
                            for (int q = 0; q < CH; ++q) {
                                const uint4 v = ldg_nc_v4(gp[q] + i * 32);
                                sad = sad4_acc(v.x, sp[q].x, sad);
                                sad = sad4_acc(v.y, sp[q].y, sad);
                                sad = sad4_acc(v.z, sp[q].z, sad);
                                sad = sad4_acc(v.w, sp[q].w, sad);
                            }
                        }
                        if (sad < ms) {                                        // std::min(d * d, minDistSq[j]) moved
                            ms = sad;
                            msnr[i] = sad | (o << 16);
                            const float d = (float)sad;
                            __stcg(mdp + i * 32, __fmul_rn(d, d));
                        }
                    }
                    const float dm = (float)ms;
                    loc32 += __float2uint_rn(__fmul_rn(dm, dm));
                }
            }
            local = (ACC)loc32;
        } else if (MDREG && !HDR) {
            // minimum distances live in registers (nseg <= 8): no loads that could alias the stores, so the
            // pixel loads of all segments are free to be issued early; a value is stored (for the leader's
            // level-2 search) only when it changed.  Addresses and validity are precomputed: about 25
            // instructions per block.  Values are integers < 2^29 and there are at most 8: a 32-bit sum is exact.
            unsigned int loc32 = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if ((vmask >> i) & 1u) {
                    unsigned int sad = 0;
                    if (i < cseg) {
#pragma unroll
                        for (int q = 0; q < CH; ++q) {
                            const uint4 v = sc[(i * CH + q) * 32];
                            sad = sad4_acc(v.x, sp[q].x, sad);
                            sad = sad4_acc(v.y, sp[q].y, sad);
                            sad = sad4_acc(v.z, sp[q].z, sad);
                            sad = sad4_acc(v.w, sp[q].w, sad);
                        }
                    } else {
#pragma unroll
                        for (int q = 0; q < CH; ++q) {
                            const uint4 v = ldg_nc_v4(gp[q] + i * 32);
                            sad = sad4_acc(v.x, sp[q].x, sad);
                            sad = sad4_acc(v.y, sp[q].y, sad);
                            sad = sad4_acc(v.z, sp[q].z, sad);
                            sad = sad4_acc(v.w, sp[q].w, sad);
                        }
                    }
                    const float d = (float)sad;                                // BlockDistanceSAD returns float(sad)
                    const float m = fminf(__fmul_rn(d, d), md[i]);             // std::min(d * d, minDistSq[j])
                    if (m != md[i]) {
                        md[i] = m;
                        __stcg(mdp + i * 32, m);
                    }
                    loc32 += __float2uint_rn(m);
                }
            }
            local = (ACC)loc32;
        } else if (TRI && MDREG && HDR) {
            // Triangle-inequality skip on the Oklab metric: sqrt(d) is a norm distance, so with s_j the seed nearest to
            // block j:  d(new, s_j) >= 4 md_j  =>  d(j, new) >= md_j  =>  the minimum is unchanged and the 256-byte
            // feature row is not read.  The factor 4.001 keeps the implication true for the fp32-computed distances
            // (64 non-negative terms: relative error < 1e-5), so values, sums and picks stay bit-identical.  The
            // distances new seed -> earlier seeds are recomputed by every CTA (one per thread, rows are L2-resident).
            // Only the seeds that are (or once were) nearest to one of THIS CTA's blocks are needed: a CTA covers a
            // contiguous stripe of the image, so that is a small part of the list (bit mask in shared memory).
            const uint32_t o = round - 1;
            float* const tri_f = reinterpret_cast<float*>(s_cache);
            uint32_t* const seedlist_s = reinterpret_cast<uint32_t*>(tri_f + K);
            uint32_t* const used_s = seedlist_s + K;                               // K / 32 words (zeroed before round 1)
            if (tid < (int)o && ((used_s[tid >> 5] >> (tid & 31)) & 1u)) {
                const float4* x = reinterpret_cast<const float4*>(feat) + (size_t)seedlist_s[tid] * 16;
                DistAcc acc;
                acc.reset();
#pragma unroll
                for (int k = 0; k < 8; ++k) acc.step(x[2 * k], x[2 * k + 1], s_seedf[2 * k], s_seedf[2 * k + 1]);
                tri_f[tid] = acc.finish();
            }
            if (tid == 0) seedlist_s[o] = seed;   // read from the next round on (barriers in between)
            __syncthreads();
            bool took = false;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const size_t j = wbase + (size_t)i * 32 + lane;
                if ((vmask >> i) & 1u) {
                    float m = md[i];
                    const bool skip = msnr[i] != 0xffffu && tri_f[msnr[i]] >= 4.001f * m;
                    if (!skip) {
                        const float d = block_dist(i, j);
                        if (d < m) {                                               // std::min(d, minDistSq[j]) moved
                            m = d;
                            md[i] = d;
                            msnr[i] = o;
                            took = true;
                            __stcg(mdp + i * 32, d);
                        }
                    }
                    local += (ACC)m;
                }
            }
            if (__any_sync(0xffffffffu, took) && lane == 0) atomicOr(&used_s[o >> 5], 1u << (o & 31));
        } else if (MDREG) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const size_t j = wbase + (size_t)i * 32 + lane;
                if ((vmask >> i) & 1u) {
                    const float m = fminf(block_dist(i, j), md[i]);               // std::min(d, minDistSq[j])
                    if (m != md[i]) {
                        md[i] = m;
                        __stcg(mdp + i * 32, m);
                    }
                    local += (ACC)m;
                }
            }
        } else {
#pragma unroll 2
            for (int i = 0; i < nseg; ++i) {
                const size_t j = wbase + (size_t)i * 32 + lane;
                if (j < wend) {
                    const float m = fminf(block_dist(i, j), __ldcg(st.min_dist + j));
                    __stcg(st.min_dist + j, m);
                    local += (ACC)m;
                }
            }
        }
        local = warp_sum(local);
        __syncwarp();  // orders every lane's min_dist stores before lane 0's release
        if (lane == 0) {
            if (HDR) {
                st_relaxed_u64(wwords + 2 * gw, (unsigned long long)__double_as_longlong((double)local));
                st_release_u64(wwords + 2 * gw + 1, (unsigned long long)round);
            } else {
                st_release_u64(wwords + gw, ((unsigned long long)round << kTagShift) | (unsigned long long)local);
            }
        }
        TICK(t_upd);
        if (leader) {
            // ---- (2) all warp partials -> S, p, owning warp
            ACC v[kPPer];
            for (;;) {  // this thread's words are requested together and re-requested until every tag matches
                unsigned long long w[kPPer], tg[kPPer];
#pragma unroll
                for (int k = 0; k < kPPer; ++k) {
                    const int idx = tid * kPPer + k;
                    w[k] = 0;
                    tg[k] = (unsigned long long)round;
                    if (idx < W) {
                        if (HDR) {
                            tg[k] = ld_relaxed_u64(wwords + 2 * idx + 1);
                        } else {
                            w[k] = ld_relaxed_u64(wwords + idx);
                            tg[k] = w[k] >> kTagShift;
                            w[k] &= kPayloadMask;
                        }
                    }
                }
                bool ok = true;
#pragma unroll
                for (int k = 0; k < kPPer; ++k) ok = ok && tg[k] == (unsigned long long)round;
                if (ok) {
                    // relaxed polls (independent, one round trip for all five) + one acquire fence: pairs with the
                    // warps' release stores, so their min_dist (and HDR value) stores are visible from here on
                    __threadfence();
#pragma unroll
                    for (int k = 0; k < kPPer; ++k) {
                        const int idx = tid * kPPer + k;
                        if (HDR) v[k] = idx < W ? (ACC)__longlong_as_double((long long)ld_relaxed_u64(wwords + 2 * idx)) : (ACC)0;
                        else v[k] = (ACC)w[k];
                    }
                    break;
                }
            }
            TICK(t_p1);
            ACC thread_tot = (ACC)0;
#pragma unroll
            for (int k = 0; k < kPPer; ++k) thread_tot += v[k];
            const ACC incl = warp_incl_scan<ACC>(thread_tot, lane);
            if (lane == 31) s_scan[wid] = incl;
            if (tid == 0) s_hit = 0xffffffffu;
            __syncthreads();
            TICK(t_p2);
            if (wid == 0) {
                const ACC wt = s_scan[lane];
                const ACC wi = warp_incl_scan<ACC>(wt, lane);
                s_scan[lane] = wi - wt;                                        // exclusive prefix of the warps
                if (lane == 31) {
                    s_scan[kPWarps] = wi;                                      // S
                    ACC target = (ACC)0;
                    if ((double)wi > 0.0) {
                        const double p = __dadd_rn(__dmul_rn(canon_u, (double)wi), 0.0);
                        // LDR: first j with prefix_j >= p  <=>  prefix_j >= ceil(p) on exact integers
                        target = HDR ? (ACC)p : (ACC)(p <= 0.0 ? 0.0 : ceil(p));
                    }
                    s_target = target;
                }
            }
            __syncthreads();
            const ACC total = s_scan[kPWarps];
            if (!((double)total > 0.0)) {  // reference :299-302: fill the rest with seed 0 and stop
                for (uint32_t rr = round + tid; rr < K; rr += kPThreads) st.seed_idx[rr] = st.seed_idx[0];
                if (tid == 0) {
                    st.status[0] = round;
                    st_relaxed_u64(pick, ((unsigned long long)round << kTagShift) | kStopBit);
                }
                return;
            }
            const ACC target = s_target;
            ACC cum = s_scan[wid] + (incl - thread_tot);
            const ACC my_before = cum;
            int my_hit = -1;
#pragma unroll
            for (int k = 0; k < kPPer; ++k) {
                const int idx = tid * kPPer + k;
                if (idx < W && my_hit < 0) {
                    if (cum + v[k] >= target) my_hit = k;
                    else cum += v[k];
                }
            }
            if (my_hit >= 0) atomicMin(&s_hit, (unsigned int)(tid * kPPer + my_hit));
            __syncthreads();
            unsigned int hw = s_hit;                                           // owning warp (global index)
            const bool take_last = hw == 0xffffffffu;                          // HDR only: rounding of the tree sum
            if (take_last) hw = (unsigned int)(W - 1);
            if ((int)hw / kPPer == tid) s_target = target - cum;               // residual inside that warp (cum stopped there)
            (void)my_before;
            __syncthreads();
            TICK(t_poll);
            if (wid == 0) {
                // ---- level 2: which block of warp hw's range
                const ACC resid = s_target;
                const size_t obase = (size_t)hw * bpw;
                const size_t oend = obase + bpw < n ? obase + bpw : n;
                ACC run = (ACC)0;
                size_t chosen = n - 1;                                         // take_last / rounding fallback
                bool found = false;
                // the warp's blocks are contiguous in memory: lane l takes eight consecutive values of each
                // 256-block chunk, so prefix order is lane order and one warp scan per chunk locates the block
                for (size_t k0 = 0; k0 < bpw && !found; k0 += 256) {
                    const size_t j0 = obase + k0 + (size_t)lane * 8;
                    float vals[8];
                    if (j0 + 8 <= oend) {  // obase, bpw multiples of 32 -> 16-byte aligned
                        const float4 a = __ldcg(reinterpret_cast<const float4*>(st.min_dist + j0));
                        const float4 b = __ldcg(reinterpret_cast<const float4*>(st.min_dist + j0) + 1);
                        vals[0] = a.x; vals[1] = a.y; vals[2] = a.z; vals[3] = a.w;
                        vals[4] = b.x; vals[5] = b.y; vals[6] = b.z; vals[7] = b.w;
                    } else {
#pragma unroll
                        for (int u = 0; u < 8; ++u) vals[u] = (j0 + u < oend && k0 + (size_t)lane * 8 + u < bpw) ? __ldcg(st.min_dist + j0 + u) : 0.0f;
                    }
                    if (k0 + (size_t)lane * 8 >= bpw) {
#pragma unroll
                        for (int u = 0; u < 8; ++u) vals[u] = 0.0f;
                    }
                    ACC mine = (ACC)0;
#pragma unroll
                    for (int u = 0; u < 8; ++u) mine += (ACC)vals[u];
                    const ACC sincl = warp_incl_scan<ACC>(mine, lane);
                    const ACC chunk = __shfl_sync(0xffffffffu, sincl, 31);
                    if (!take_last && run + chunk >= resid) {
                        ACC cum2 = run + (sincl - mine);
                        int u_hit = -1;
#pragma unroll
                        for (int u = 0; u < 8; ++u) {
                            cum2 += (ACC)vals[u];
                            if (u_hit < 0 && j0 + u < oend && cum2 >= resid) u_hit = u;
                        }
                        const unsigned hit = __ballot_sync(0xffffffffu, u_hit >= 0);
                        if (hit) {
                            const int f = __ffs(hit) - 1;
                            chosen = obase + k0 + (size_t)f * 8 + (size_t)__shfl_sync(0xffffffffu, u_hit, f);
                            found = true;
                        }
                    }
                    run += chunk;
                }
                if (lane == 0) {
                    st.seed_idx[round] = (uint32_t)chosen;
                    st.status[1] = round;                                      // rounds that drew a sample
                    st_relaxed_u64(pick, ((unsigned long long)round << kTagShift) | (unsigned long long)chosen);
                }
            }
            TICK(t_l2);
        }
        // ---- (3) everyone: wait for the pick
        if (tid == 0) {
            unsigned long long w;
            do { w = ld_relaxed_u64(pick); } while ((w >> kTagShift) != (unsigned long long)round);
            s_pick = w;
        }
        __syncthreads();
        if (s_pick & kStopBit) return;
        seed = (uint32_t)(s_pick & 0xffffffffull);
        TICK(t_pick);
    }
#ifdef YUPVQ_SEED_TIMING
    if (tid == 0 && (blockIdx.x == 0 || blockIdx.x == 77))
        printf("cta %d: cycles/round update %lld, poll %lld, sync1 %lld, scan %lld, level2 %lld, pick %lld\n", blockIdx.x, t_upd / (K - 1),
               t_p1 / (K - 1), t_p2 / (K - 1), t_poll / (K - 1), t_l2 / (K - 1), t_pick / (K - 1));
#endif
}

// ---- K2p: K-Means++ seeding of BC1 blocks from their palette form ---------------------------------------------------
// Same rounds, same integer sums, same picks as seed_persistent_kernel<false, 3, true> (reference :287-313), for the
// one input the headline path has: BC1 blocks with 3-channel output.  Two things change.
//  (1) Data.  A BC1 block is 20 bytes in palette form (convert.cu bc1_palette_planes_kernel: the four palette entries
//      of R, G, B and sixteen selector nibbles), so ALL blocks of a CTA are resident in shared memory (160 KiB) and
//      the minimum distances live in registers: a round reads and writes no global memory except the exchange words.
//      A lane rebuilds the 48 pixel bytes channel by channel with 12 PRMTs and feeds 12 VABSDIFF4 against the seed
//      rebuilt the same way; SAD adds the same 48 |a - b| in another order.
//  (2) The chain of a round.  The generic kernel publishes 32 warp words per CTA with a release (its min_dist stores must
//      be visible to the leader's level-2 search), the leader scans 4736 words with three CTA barriers, reads the
//      owner warp's min_dist from L2 and publishes the pick; every CTA then fetches the seed's pixels: five dependent
//      L2 round trips.  Here the search goes to the data instead: each CTA publishes ONE tagged word (its total), one
//      warp of the leader scans the <= 160 words and posts the residual target in the owning CTA's mailbox; that
//      CTA finds the block in its own registers (warp totals are still in shared memory) and publishes the pick
//      TOGETHER with the seed's 20 bytes taken from its shared memory.  Three round trips, every message a
//      self-validating tagged 64-bit word (16-bit round tag + 48-bit payload): no fences, no min_dist array.
constexpr int kQMailbox = 256, kQMailStride = 16, kQPick = 4096;  // u64 offsets in SeedState::partial (CTA words at 0)
constexpr unsigned int kEvtPick = 1, kEvtOwner = 2;

__global__ void __launch_bounds__(kPThreads, 1)
seed_palette_kernel(SeedState st, const uint32_t* __restrict__ pal5, size_t pstride, size_t n, size_t bpw, uint32_t K) {
    extern __shared__ uint4 s_cache[];                 // [warp][8 segments][5 words][32 lanes]
    __shared__ unsigned long long s_w[kPWarps];        // warp totals of the round
    __shared__ unsigned long long s_pickw[5];
    __shared__ unsigned long long s_resid;
    __shared__ unsigned int s_evt;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int G = gridDim.x;
    const size_t gw = (size_t)blockIdx.x * kPWarps + wid;
    const size_t wbase = gw * bpw;
    const size_t wend = wbase + bpw < n ? wbase + bpw : n;  // may be <= wbase: empty warp
    const int nseg = (int)(bpw / 32);
    unsigned long long* const cta_words = st.partial;
    unsigned long long* const mailbox = st.partial + kQMailbox + (size_t)blockIdx.x * kQMailStride;
    unsigned long long* const pickw = st.partial + kQPick;
    const bool leader = blockIdx.x == 0;

    uint32_t* const pc = reinterpret_cast<uint32_t*>(s_cache) + (size_t)wid * 8 * 5 * 32 + lane;
    for (int i = 0; i < nseg; ++i) {
        const size_t j = wbase + (size_t)i * 32 + lane;
#pragma unroll
        for (int q = 0; q < 5; ++q) pc[(i * 5 + q) * 32] = j < wend ? __ldg(pal5 + q * pstride + j) : 0u;
    }
    __syncwarp();
    unsigned int vmask = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i)
        if (i < nseg && wbase + (size_t)i * 32 + lane < wend) vmask |= 1u << i;
    // minDistSq[j] = float(sad) * float(sad) is monotone in sad, so a lane keeps the minimum SAD per block and the
    // running sum of the (integer-valued) squares of its blocks, which only moves when a minimum moves
    auto sq_of = [](uint32_t sad) -> uint32_t {
        const float d = (float)sad;                                            // BlockDistanceSAD returns float(sad)
        return __float2uint_rn(__fmul_rn(d, d));                               // d * d, an integer-valued float < 2^28
    };
    uint32_t ms[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) ms[i] = 0xffffffffu;
    uint32_t lsum = 0;  // at most 8 values < 2^28
    uint32_t c[5];
    {
        const uint32_t seed = st.seed_idx[0];
#pragma unroll
        for (int q = 0; q < 5; ++q) c[q] = __ldg(pal5 + q * pstride + seed);
    }
#ifdef YUPVQ_SEED_TIMING
    long long q_upd = 0, q_bar = 0, q_lead = 0, q_poll = 0, q_own = 0, t0 = 0, t1;
#endif
    for (uint32_t round = 1; round < K; ++round) {
#ifdef YUPVQ_SEED_TIMING
        t0 = clock64();
#endif
        const unsigned long long tag = (unsigned long long)round << kTagShift;
        const double canon_u = (leader && wid == 0) ? st.canon[round - 1] : 0.0;  // fetched early: off the critical path
        // ---- (1) the new seed's pixels, channel-planar: sp[q] = channel q of texels 0-3, 4-7, 8-11, 12-15
        uint4 sp[3];
        {
            const uint32_t s1 = c[3] >> 16, s3 = c[4] >> 16;
#pragma unroll
            for (int q = 0; q < 3; ++q)
                sp[q] = make_uint4(prmt_lo(c[q], c[3]), prmt_lo(c[q], s1), prmt_lo(c[q], c[4]), prmt_lo(c[q], s3));
        }
        // ---- (2) min-distance update, all operands in shared memory and registers
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if ((vmask >> i) & 1u) {
                const uint32_t sa = pc[(i * 5 + 3) * 32], sb = pc[(i * 5 + 4) * 32];
                const uint32_t s1 = sa >> 16, s3 = sb >> 16;
                uint32_t part[3];  // one accumulate chain per channel
#pragma unroll
                for (int q = 0; q < 3; ++q) {
                    const uint32_t pw = pc[(i * 5 + q) * 32];
                    part[q] = sad4_acc(prmt_lo(pw, sa), sp[q].x, 0u);
                    part[q] = sad4_acc(prmt_lo(pw, s1), sp[q].y, part[q]);
                    part[q] = sad4_acc(prmt_lo(pw, sb), sp[q].z, part[q]);
                    part[q] = sad4_acc(prmt_lo(pw, s3), sp[q].w, part[q]);
                }
                const unsigned int sad = part[0] + part[1] + part[2];
                if (sad < ms[i]) {                                             // std::min(d * d, minDistSq[j]) moved
                    lsum += sq_of(sad) - (ms[i] == 0xffffffffu ? 0u : sq_of(ms[i]));
                    ms[i] = sad;
                }
            }
        }
        // warp total < 2^36: two hardware reductions over the 16-bit halves
        const unsigned long long wt = ((unsigned long long)__reduce_add_sync(0xffffffffu, lsum >> 16) << 16) +
                                      (unsigned long long)__reduce_add_sync(0xffffffffu, lsum & 0xffffu);
        if (lane == 0) s_w[wid] = wt;
        TICK(q_upd);
        __syncthreads();
        TICK(q_bar);
        if (wid == 0) {
            const unsigned long long wv = s_w[lane];   // < 2^36; CTA total < 2^41
            const unsigned long long cta_total = ((unsigned long long)__reduce_add_sync(0xffffffffu, (unsigned int)(wv >> 20)) << 20) +
                                                 (unsigned long long)__reduce_add_sync(0xffffffffu, (unsigned int)(wv & 0xfffffu));
            if (lane == 0) st_relaxed_u64(cta_words + blockIdx.x, tag | cta_total);
            if (leader) {
                // ---- (3) leader warp: all CTA totals -> S, target, owning CTA, residual
                unsigned long long v[kPPer];
                for (;;) {
                    bool ok = true;
#pragma unroll
                    for (int k = 0; k < kPPer; ++k) {
                        const int idx = lane * kPPer + k;
                        unsigned long long w = tag;
                        if (idx < G) w = ld_relaxed_u64(cta_words + idx);
                        ok = ok && (w >> kTagShift) == (unsigned long long)round;
                        v[k] = w & kPayloadMask;
                    }
                    if (__all_sync(0xffffffffu, ok)) break;
                }
                unsigned long long thread_tot = 0;
#pragma unroll
                for (int k = 0; k < kPPer; ++k) thread_tot += v[k];
                const unsigned long long incl = warp_incl_scan<unsigned long long>(thread_tot, lane);
                const unsigned long long S = __shfl_sync(0xffffffffu, incl, 31);
                if (S == 0) {  // reference :299-302: fill the rest with seed 0 and stop
                    for (uint32_t rr = round + lane; rr < K; rr += 32) st.seed_idx[rr] = st.seed_idx[0];
                    if (lane == 0) st.status[0] = round;
                    if (lane < 5) st_relaxed_u64(pickw + lane, tag | kStopBit);
                } else {
                    const double p = __dadd_rn(__dmul_rn(canon_u, (double)S), 0.0);
                    // first j with prefix_j >= p  <=>  prefix_j >= ceil(p) on exact integers
                    const unsigned long long target = (unsigned long long)(p <= 0.0 ? 0.0 : ceil(p));
                    unsigned long long cum = incl - thread_tot;
                    int hit = -1;
#pragma unroll
                    for (int k = 0; k < kPPer; ++k) {
                        if (lane * kPPer + k < G && hit < 0) {
                            if (cum + v[k] >= target) hit = k;
                            else cum += v[k];
                        }
                    }
                    const unsigned int hb = __ballot_sync(0xffffffffu, hit >= 0);
                    int f = hb ? __ffs(hb) - 1 : (G - 1) / kPPer;                // no hit: cannot happen on exact integers
                    int hk = __shfl_sync(0xffffffffu, hit, f);
                    if (hk < 0) hk = (G - 1) % kPPer;
                    const unsigned long long before = __shfl_sync(0xffffffffu, cum, f);
                    const unsigned long long resid = target > before ? target - before : 0ull;
                    if (lane == 0) {
                        st.status[1] = round;                                    // rounds that drew a sample
                        st_relaxed_u64(st.partial + kQMailbox + (size_t)(f * kPPer + hk) * kQMailStride, tag | resid);
                    }
                }
            }
        }
        TICK(q_lead);
        // ---- (4) wait for the pick; the CTA whose mailbox is posted first finds the block and publishes it
        bool owner_done = false;
        for (;;) {
            if (wid == 0) {
                unsigned long long w = 0;
                unsigned int b;
                for (;;) {
                    if (lane < 5) w = ld_relaxed_u64(pickw + lane);
                    else if (lane == 5 && !owner_done) w = ld_relaxed_u64(mailbox);
                    const bool ok = lane < 6 && (w >> kTagShift) == (unsigned long long)round;
                    b = __ballot_sync(0xffffffffu, ok);
                    if ((b & 0x1fu) == 0x1fu || (b & 0x20u)) break;
                }
                if ((b & 0x1fu) == 0x1fu) {
                    if (lane < 5) s_pickw[lane] = w;
                    if (lane == 0) s_evt = kEvtPick;
                } else {
                    if (lane == 5) s_resid = w & kPayloadMask;
                    if (lane == 0) s_evt = kEvtOwner;
                    owner_done = true;
                }
            }
            __syncthreads();
            TICK(q_poll);
            if (s_evt == kEvtPick) break;
            {   // owner: first block (in block order) whose inclusive prefix inside this CTA reaches the residual
                const unsigned long long resid = s_resid;
                const unsigned long long wtot = s_w[lane];
                const unsigned long long wincl = warp_incl_scan<unsigned long long>(wtot, lane);
                const unsigned int wb = __ballot_sync(0xffffffffu, wincl >= resid);
                const int ow = wb ? __ffs(wb) - 1 : kPWarps - 1;
                const unsigned long long before_w = __shfl_sync(0xffffffffu, wincl - wtot, ow);
                if (wid == ow) {
                    const unsigned long long r = resid - before_w;
                    // segment totals by hardware reduction (independent, pipelined), then ONE scan inside the segment
                    // that holds the residual
                    uint32_t v32[8];
                    unsigned long long run = 0;
                    int si = -1;
#pragma unroll
                    for (int i = 0; i < 8; ++i) v32[i] = ((vmask >> i) & 1u) ? sq_of(ms[i]) : 0u;
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const unsigned long long tot = ((unsigned long long)__reduce_add_sync(0xffffffffu, v32[i] >> 16) << 16) +
                                                       (unsigned long long)__reduce_add_sync(0xffffffffu, v32[i] & 0xffffu);
                        if (si < 0) {
                            if (run + tot >= r) si = i;
                            else run += tot;
                        }
                    }
                    bool found = false;
                    if (si >= 0) {
                        uint32_t v = 0;
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            if (i == si) v = v32[i];
                        const unsigned long long incl = warp_incl_scan<unsigned long long>((unsigned long long)v, lane);
                        const unsigned int hb = __ballot_sync(0xffffffffu, run + incl >= r);
                        if (hb) {
                            found = true;
                            if (lane == __ffs(hb) - 1) {
                                const size_t chosen = wbase + (size_t)si * 32 + lane;
                                uint32_t e[5];
#pragma unroll
                                for (int q = 0; q < 5; ++q) e[q] = pc[(si * 5 + q) * 32];
                                st.seed_idx[round] = (uint32_t)chosen;
                                st_relaxed_u64(pickw + 0, tag | (unsigned long long)chosen);
                                st_relaxed_u64(pickw + 1, tag | (unsigned long long)e[0] | ((unsigned long long)(e[1] & 0xffffu) << 32));
                                st_relaxed_u64(pickw + 2, tag | (unsigned long long)(e[1] >> 16) | ((unsigned long long)e[2] << 16));
                                st_relaxed_u64(pickw + 3, tag | (unsigned long long)e[3] | ((unsigned long long)(e[4] & 0xffffu) << 32));
                                st_relaxed_u64(pickw + 4, tag | (unsigned long long)(e[4] >> 16));
                            }
                        }
                    }
                    if (!found && lane == 0) {  // cannot happen (r <= warp total); never leave the grid waiting
                        st.seed_idx[round] = (uint32_t)(n - 1);
#pragma unroll
                        for (int q = 0; q < 5; ++q) st_relaxed_u64(pickw + q, tag | kStopBit);
                        st.status[0] = round;
                    }
                }
            }
            __syncthreads();
            TICK(q_own);
        }
        const unsigned long long p0 = s_pickw[0], p1 = s_pickw[1], p2 = s_pickw[2], p3 = s_pickw[3], p4 = s_pickw[4];
        if (p0 & kStopBit) return;
        c[0] = (uint32_t)(p1 & 0xffffffffull);
        c[1] = (uint32_t)((p1 >> 32) & 0xffffull) | ((uint32_t)(p2 & 0xffffull) << 16);
        c[2] = (uint32_t)((p2 >> 16) & 0xffffffffull);
        c[3] = (uint32_t)(p3 & 0xffffffffull);
        c[4] = (uint32_t)((p3 >> 32) & 0xffffull) | ((uint32_t)(p4 & 0xffffull) << 16);
    }
#ifdef YUPVQ_SEED_TIMING
    if (tid == 0 && (blockIdx.x == 0 || blockIdx.x == 77))
        printf("cta %d: cycles/round update %lld, barrier %lld, leader %lld, poll %lld, owner %lld\n", blockIdx.x, q_upd / (K - 1),
               q_bar / (K - 1), q_lead / (K - 1), q_poll / (K - 1), q_own / (K - 1));
#endif
}

inline size_t chunk_for(size_t n, int grid) {
    size_t c = (n + grid - 1) / grid;
    return (c + kThreads - 1) / kThreads * kThreads;
}

}  // namespace

// bytes of SeedState::partial: per-round partial sums of the launch-per-round kernels, or the tagged
// exchange words of the persistent kernel ([2 * kPMaxGrid] slots + the pick word)
size_t seed_partial_bytes(int grid) {
    const size_t a = (size_t)(grid > 0 ? grid : 0) * 8, b = (size_t)(2 * kPMaxGrid * kPWarps + 8) * 8;
    return a > b ? a : b;
}

size_t seed_grid_for(size_t n) {
    size_t g = (n + kThreads - 1) / kThreads;
    if (g > (size_t)kMaxGrid) g = kMaxGrid;
    if (g < 1) g = 1;
    // every CTA must own a non-empty chunk so that the ticket count equals the launched grid
    const size_t c = chunk_for(n, (int)g);
    g = (n + c - 1) / c;
    return g;
}

cudaError_t launch_seed_init(const SeedState& st, size_t n, uint32_t first_idx, cudaStream_t s) {
    cudaError_t e = cudaMemsetAsync(st.partial, 0, seed_partial_bytes(0), s);  // tagged words of the persistent kernel
    if (e != cudaSuccess) return e;
    seed_init_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(st.min_dist, n, st.counter, st.status, st.seed_idx,
                                                                 first_idx);
    return cudaGetLastError();
}

cudaError_t launch_seed_round_ldr(const SeedState& st, const uint4* planes, size_t plane_stride, size_t n, int ch,
                                  uint32_t round, uint32_t K, cudaStream_t s) {
    (void)K;
    const size_t chunk = chunk_for(n, st.grid);
    switch (ch) {
        case 1: seed_round_ldr_kernel<1><<<st.grid, kThreads, 0, s>>>(st, planes, plane_stride, n, chunk, round); break;
        case 2: seed_round_ldr_kernel<2><<<st.grid, kThreads, 0, s>>>(st, planes, plane_stride, n, chunk, round); break;
        case 3: seed_round_ldr_kernel<3><<<st.grid, kThreads, 0, s>>>(st, planes, plane_stride, n, chunk, round); break;
        case 4: seed_round_ldr_kernel<4><<<st.grid, kThreads, 0, s>>>(st, planes, plane_stride, n, chunk, round); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

cudaError_t launch_seed_persistent(const SeedState& st, const uint4* planes, size_t plane_stride, const float* feat, size_t n,
                                   int ch, bool hdr, uint32_t K, cudaStream_t s, const uint32_t* pal5, size_t pal_stride) {
    if (K <= 1) return cudaSuccess;
    int sms = 0, coop = 0;  // per device: queried on every call (a process may hold contexts on several GPUs)
    {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    }
    if (!coop || sms <= 0) return cudaErrorNotSupported;
    size_t g = (n + kPThreads - 1) / kPThreads;
    if (g > (size_t)sms) g = sms;
    if (g > (size_t)kPMaxGrid) g = kPMaxGrid;
    if (g < 1) g = 1;
    size_t bpw = (n + g * kPWarps - 1) / (g * kPWarps);
    bpw = (bpw + 31) / 32 * 32;
    if (!hdr && bpw * kPWarps >= ((size_t)1 << 19)) return cudaErrorNotSupported;  // CTA partial must fit 48 bits
    const bool mdreg = bpw / 32 <= 8;
    // Measured on config 2 (B200): the skip makes the seeding SLOWER (4.41 ms against 3.84 ms) -- the extra barrier and
    // table per round cost more than the skipped loads save, because a warp only saves a load when all 32 of its
    // (noisy, 48-dimensional) blocks pass the bound.  Kept opt-in for smoother inputs; results are identical either way.
    // The Oklab (HDR) variant is on by default: its rows are 256 bytes (the round is HBM-bound) and neighbouring
    // blocks of real images are similar, so whole warps skip (YUPVQ_SEED_NO_TRI=1 turns it off for A/B).
    const bool want_tri = debug_opts().seed_tri;
    const bool no_tri = debug_opts().seed_no_tri;
    const bool tri = !hdr && mdreg && K <= (uint32_t)kPThreads && want_tri;
    const bool tri_hdr = hdr && mdreg && K <= (uint32_t)kPThreads && !no_tri;
    const void* fn = nullptr;
    if (hdr) fn = mdreg ? (const void*)seed_persistent_kernel<true, 1, true> : (const void*)seed_persistent_kernel<true, 1, false>;
    else if (ch == 1) fn = mdreg ? (const void*)seed_persistent_kernel<false, 1, true> : (const void*)seed_persistent_kernel<false, 1, false>;
    else if (ch == 2) fn = mdreg ? (const void*)seed_persistent_kernel<false, 2, true> : (const void*)seed_persistent_kernel<false, 2, false>;
    else if (ch == 3) fn = mdreg ? (const void*)seed_persistent_kernel<false, 3, true> : (const void*)seed_persistent_kernel<false, 3, false>;
    else if (ch == 4) fn = mdreg ? (const void*)seed_persistent_kernel<false, 4, true> : (const void*)seed_persistent_kernel<false, 4, false>;
    else return cudaErrorInvalidValue;
    if (tri_hdr) fn = (const void*)seed_persistent_kernel<true, 1, true, true>;
    if (tri) {
        if (ch == 1) fn = (const void*)seed_persistent_kernel<false, 1, true, true>;
        else if (ch == 2) fn = (const void*)seed_persistent_kernel<false, 2, true, true>;
        else if (ch == 3) fn = (const void*)seed_persistent_kernel<false, 3, true, true>;
        else fn = (const void*)seed_persistent_kernel<false, 4, true, true>;
    }
    // BC1 input in palette form: its own kernel (A/B: YUPVQ_SEED_NO_PAL=1 keeps the pixel-plane kernel)
    const bool no_pal = debug_opts().seed_no_pal;
    if (pal5 != nullptr && !hdr && ch == 3 && mdreg && !tri && !no_pal && K <= 0xffffu) {
        const size_t smem_pal = (size_t)kPWarps * 8 * 5 * 32 * 4;  // 160 KiB
        static std::atomic<uint64_t> attr_pal{0};
        if (cudaError_t e = once_per_device(attr_pal, [&] {
                return cudaFuncSetAttribute(seed_palette_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_pal);
            }); e != cudaSuccess)
            return e;
        SeedState st_pal = st;
        void* pargs[] = {&st_pal, (void*)&pal5, &pal_stride, &n, &bpw, &K};
        return cudaLaunchCooperativeKernel((const void*)seed_palette_kernel, dim3((unsigned)g), dim3(kPThreads), pargs, smem_pal, s);
    }
    // shared-memory resident segments per warp: as many as fit in ~200 KB (after the seed table of the TRI variant)
    int cseg = 0;
    size_t smem = 0;
    if (!hdr) {
        const size_t seg_bytes = (size_t)kPWarps * 32 * ch * 16;
        const size_t tri_bytes = tri ? (size_t)K * ch * 16 + (size_t)K * 4 : 0;
        cseg = (int)((200u * 1024u - tri_bytes) / seg_bytes);
        if ((size_t)cseg > bpw / 32) cseg = (int)(bpw / 32);
        if (debug_opts().seed_no_smem) cseg = 0;
        smem = (size_t)cseg * seg_bytes + tri_bytes;
        static std::atomic<uint64_t> attr[24];  // dynamic + static shared memory may exceed the 48 KB default: always opt in
        const int ai = ch * 2 + (mdreg ? 1 : 0) + (tri ? 12 : 0);
        if (smem > 0)
            if (cudaError_t e = once_per_device(attr[ai], [&] {
                    return cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(200u * 1024u));
                }); e != cudaSuccess)
                return e;
    }
    if (tri_hdr) {  // distances to the earlier seeds + their block indices
        smem = (size_t)K * 8 + ((size_t)K + 31) / 32 * 4;
        static std::atomic<uint64_t> attr_hdr{0};
        if (smem > 40 * 1024)
            if (cudaError_t e = once_per_device(attr_hdr, [&] {
                    return cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(64u * 1024u));
                }); e != cudaSuccess)
                return e;
    }
    SeedState st_copy = st;
    void* args[] = {&st_copy, (void*)&planes, &plane_stride, (void*)&feat, &n, &bpw, &cseg, &K};
    return cudaLaunchCooperativeKernel(fn, dim3((unsigned)g), dim3(kPThreads), args, smem, s);
}

cudaError_t launch_seed_round_hdr(const SeedState& st, const float* feat, size_t n, uint32_t round, uint32_t K,
                                  cudaStream_t s) {
    (void)K;
    const size_t chunk = chunk_for(n, st.grid);
    seed_round_hdr_kernel<<<st.grid, kThreads, 0, s>>>(st, feat, n, chunk, round);
    return cudaGetLastError();
}

}  // namespace yupvq
