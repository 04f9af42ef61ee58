// K4 / K5: centroid sum/count update, normalisation and empty-cluster reseed.
// Reference: src/vq_encoder.cpp:343-375 (LDR) and :574-605 (HDR).
//
// Sums are accumulated in 64-bit fixed point (value * 2^shift, shift chosen from N and max|x| so
// the total cannot overflow).  Integer addition is associative, so the result does not depend on
// thread/CTA/GPU order -- 1, 2, 4 and 8 GPUs produce identical centroids -- and it is rounded to
// fp32 exactly once, which is closer to the true mean than the reference's thread-chunked fp32
// accumulation (whose own result depends on numThreads).  Normalisation follows the reference's
// two roundings: c = float(sum) * (1.0f / count).
// Accumulator layout (int64 words): [K*64 sums | K counts | 1 changed flag].
// Roofline: HBM-bound, 260 B read per block per pass.
#include <cfloat>

#include "common.cuh"
#include "host_util.h"
#include "kernels.h"

namespace yupvq {

namespace {

// v1: one warp per block, 64 RED.ADD.64 per block into the accumulator in L2.
__global__ void __launch_bounds__(256) accumulate_kernel(const float* __restrict__ feat, size_t n,
                                                         const uint32_t* __restrict__ assign, uint32_t K, float scale,
                                                         long long* __restrict__ acc, const uint32_t* __restrict__ flags,
                                                         const uint32_t* __restrict__ stop, const uint32_t* __restrict__ dctl) {
    if (stop && __ldg(stop)) return;
    if (dctl && dctl[33] == 0u) return;  // this pass is a delta update
    const int lane = threadIdx.x & 31;
    const size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const size_t n_warps = ((size_t)gridDim.x * blockDim.x) >> 5;
    unsigned long long* sums = reinterpret_cast<unsigned long long*>(acc);
    unsigned long long* counts = sums + (size_t)K * 64;
    for (size_t i = warp; i < n; i += n_warps) {
        const uint32_t c = assign[i];
        const float v0 = feat[i * 64 + lane], v1 = feat[i * 64 + 32 + lane];
        atomicAdd(sums + (size_t)c * 64 + lane, (unsigned long long)__float2ll_rn(__fmul_rn(v0, scale)));
        atomicAdd(sums + (size_t)c * 64 + 32 + lane, (unsigned long long)__float2ll_rn(__fmul_rn(v1, scale)));
        if (lane == 0) atomicAdd(counts + c, 1ull);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && flags) atomicAdd(sums + (size_t)K * 65, (unsigned long long)flags[0]);
}

// v2: sort-and-own.  A CTA (two per SM) takes chunks of kAccChunk consecutive blocks, counting-sorts the chunk's
// assignments in shared memory (histogram, scan, scatter of 16-bit local ids), then its warps claim
// centroids and sum each centroid's member rows in registers: a row is one coalesced 256-byte read
// (float4 per lane of a half-warp), the fixed-point conversion happens once per element, and only one 64-bit RED
// per (chunk, centroid, column) reaches L2 -- 8x fewer than one per (block, column) at K = 512.  No shared
// memory atomics on the data path; integer sums keep the result order-independent.
constexpr int kAccThreads = 512;
constexpr int kAccChunk = 4096;
constexpr int kAccMaxK = 4096;

__device__ __forceinline__ void
accumulate_sorted_body(const float* __restrict__ feat_all, size_t n, const uint32_t* __restrict__ assign_all, uint32_t K, float scale1,
                       long long* __restrict__ acc_all, const uint32_t* __restrict__ flags, const uint32_t* __restrict__ stop,
                       const AccBatch& bt, const uint32_t* __restrict__ dctl) {
    // Lock-step batch (bt.n_textures > 1): texture t owns rows [t * n, (t + 1) * n) of feat / assign, the accumulator
    // at acc + t * acc_stride, the scale bt.scales[t] and the flag / stop words at the given strides; chunks are
    // numbered texture-major, so one launch covers the whole batch.
    extern __shared__ uint32_t sm_acc[];
    if (bt.n_textures <= 1 && stop && __ldg(stop)) return;
    if (dctl && dctl[33] == 0u) return;  // this pass is a delta update
    uint32_t* hist = sm_acc;                                      // K
    uint32_t* offs = sm_acc + K;                                  // K
    uint16_t* sorted = reinterpret_cast<uint16_t*>(sm_acc + 2 * K);  // kAccChunk
    __shared__ uint32_t warp_tot[kAccThreads / 32 + 1];
    __shared__ uint32_t next_c;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const size_t n_chunks = (n + kAccChunk - 1) / kAccChunk;     // per texture
    constexpr int kPer = kAccChunk / kAccThreads;                 // blocks per thread per chunk
    for (size_t gchunk = blockIdx.x; gchunk < n_chunks * bt.n_textures; gchunk += gridDim.x) {
        const uint32_t tex = (uint32_t)(gchunk / n_chunks);
        if (bt.n_textures > 1 && stop && __ldg(stop + (size_t)tex * bt.stop_stride)) continue;  // uniform across the CTA
        const size_t chunk = gchunk - (size_t)tex * n_chunks;
        const float* feat = feat_all + (size_t)tex * n * 64;
        const uint32_t* assign = assign_all + (size_t)tex * n;
        const float scale = bt.scales ? __ldg(bt.scales + tex) : scale1;
        unsigned long long* sums = reinterpret_cast<unsigned long long*>(acc_all) + (size_t)tex * bt.acc_stride;
        unsigned long long* counts = sums + (size_t)K * 64;
        const size_t first = chunk * kAccChunk;
        const uint32_t len = (uint32_t)((n - first) < (size_t)kAccChunk ? (n - first) : (size_t)kAccChunk);
        for (uint32_t c = tid; c < K; c += kAccThreads) hist[c] = 0;
        if (tid == 0) next_c = 0;
        __syncthreads();
        uint32_t a[kPer], rank[kPer];
#pragma unroll
        for (int k = 0; k < kPer; ++k) {
            const uint32_t i = (uint32_t)k * kAccThreads + tid;
            a[k] = i < len ? assign[first + i] : 0xffffffffu;
            rank[k] = a[k] < K ? atomicAdd(&hist[a[k]], 1u) : 0u;
        }
        __syncthreads();
        // exclusive scan of hist -> offs (each thread owns a contiguous run of ceil(K / threads) bins)
        {
            const uint32_t per = (K + kAccThreads - 1) / kAccThreads;
            const uint32_t b0 = tid * per;
            uint32_t local = 0;
            for (uint32_t c = b0; c < b0 + per && c < K; ++c) local += hist[c];
            uint32_t incl = local;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            if (lane == 31) warp_tot[wid] = incl;
            __syncthreads();
            if (wid == 0) {
                uint32_t w = lane < kAccThreads / 32 ? warp_tot[lane] : 0u, wi = w;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
                    if (lane >= o) wi += t;
                }
                if (lane < kAccThreads / 32) warp_tot[lane] = wi - w;
            }
            __syncthreads();
            uint32_t run = warp_tot[wid] + incl - local;
            for (uint32_t c = b0; c < b0 + per && c < K; ++c) {
                offs[c] = run;
                run += hist[c];
            }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kPer; ++k)
            if (a[k] < K) sorted[offs[a[k]] + rank[k]] = (uint16_t)((uint32_t)k * kAccThreads + tid);
        __syncthreads();
        // warps claim centroids; a half-warp reads one member row (float4 per lane, 256 B coalesced), so a
        // step of 8 loads per lane covers 16 rows; sums live in four 64-bit registers per lane
        const float4* rows = reinterpret_cast<const float4*>(feat) + first * 16 + (lane & 15);
        const uint32_t hw = (uint32_t)lane >> 4;
        for (;;) {
            uint32_t c = 0;
            if (lane == 0) c = atomicAdd(&next_c, 1u);
            c = __shfl_sync(0xffffffffu, c, 0);
            if (c >= K) break;
            const uint32_t m = hist[c];
            if (m == 0) continue;
            const uint16_t* ids = sorted + offs[c];
            long long s0 = 0, s1 = 0, s2 = 0, s3 = 0;
            for (uint32_t j = 0; j < m; j += 16) {
                float4 v[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const uint32_t r = j + 2 * q + hw;
                    v[q] = r < m ? __ldg(rows + (size_t)ids[r] * 16) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    if (j + 2 * q + hw < m) {
                        s0 += __float2ll_rn(__fmul_rn(v[q].x, scale));
                        s1 += __float2ll_rn(__fmul_rn(v[q].y, scale));
                        s2 += __float2ll_rn(__fmul_rn(v[q].z, scale));
                        s3 += __float2ll_rn(__fmul_rn(v[q].w, scale));
                    }
                }
            }
            s0 += __shfl_xor_sync(0xffffffffu, s0, 16);
            s1 += __shfl_xor_sync(0xffffffffu, s1, 16);
            s2 += __shfl_xor_sync(0xffffffffu, s2, 16);
            s3 += __shfl_xor_sync(0xffffffffu, s3, 16);
            if (lane < 16) {
                unsigned long long* dst = sums + (size_t)c * 64 + 4 * lane;
                atomicAdd(dst + 0, (unsigned long long)s0);
                atomicAdd(dst + 1, (unsigned long long)s1);
                atomicAdd(dst + 2, (unsigned long long)s2);
                atomicAdd(dst + 3, (unsigned long long)s3);
            }
            if (lane == 0) atomicAdd(counts + c, (unsigned long long)m);
        }
        __syncthreads();
    }
    if (blockIdx.x == 0 && flags) {
        for (uint32_t tex = tid; tex < bt.n_textures; tex += kAccThreads) {
            if (bt.n_textures > 1 && stop && __ldg(stop + (size_t)tex * bt.stop_stride)) continue;
            unsigned long long* sums = reinterpret_cast<unsigned long long*>(acc_all) + (size_t)tex * bt.acc_stride;
            atomicAdd(sums + (size_t)K * 65, (unsigned long long)flags[(size_t)tex * bt.flags_stride]);
        }
    }
}

__global__ void __launch_bounds__(kAccThreads, 2)
accumulate_sorted_kernel(const float* __restrict__ feat_all, size_t n, const uint32_t* __restrict__ assign_all, uint32_t K, float scale1,
                         long long* __restrict__ acc_all, const uint32_t* __restrict__ flags, const uint32_t* __restrict__ stop,
                         AccBatch bt, const uint32_t* __restrict__ dctl) {
    accumulate_sorted_body(feat_all, n, assign_all, K, scale1, acc_all, flags, stop, bt, dctl);
}

// ---- delta update ----------------------------------------------------------------------------------------------
// After the first passes only a few percent of the blocks change centroid.  Because the sums are integers,
// sums[new] += x, sums[old] -= x over the changed rows gives exactly the accumulator a full recompute would give,
// while reading only those rows.  The assignment kernels append (row, previous centroid) for every row they move.
constexpr int kDeltaThreads = 512;
constexpr int kDeltaRows = 2048;           // listed rows per chunk -> 2 * kDeltaRows signed entries
constexpr int kDeltaMaxK = 4096;

// one CTA: decide this pass' mode, zero what has to be zeroed, hand the row count over and rearm the counter
// (lock-step batch: n_textures accumulators acc_stride words apart share one list and one mode)
__global__ void __launch_bounds__(1024) acc_prepare_kernel(long long* __restrict__ acc, uint32_t K, uint32_t full_above,
                                                           uint32_t force_full, uint32_t* __restrict__ dctl,
                                                           const uint32_t* __restrict__ stop, uint32_t n_textures, size_t acc_stride) {
    if (stop && __ldg(stop)) return;
    const uint32_t n_chg = dctl[0];
    const bool full = force_full || n_chg > full_above;
    __syncthreads();
    if (threadIdx.x == 0) {
        dctl[32] = n_chg;
        dctl[33] = full ? 1u : 0u;
        dctl[0] = 0u;
    }
    for (uint32_t t = threadIdx.x; t < n_textures; t += blockDim.x) acc[(size_t)t * acc_stride + (size_t)K * 65] = 0;  // changed word: per pass
    if (full)
        for (uint32_t t = 0; t < n_textures; ++t)
            for (size_t i = threadIdx.x; i < (size_t)K * 65; i += blockDim.x) acc[(size_t)t * acc_stride + i] = 0;
}

__device__ __forceinline__ void
accumulate_delta_body(const float* __restrict__ feat, const uint32_t* __restrict__ assign, uint32_t K1, float scale1,
                      long long* __restrict__ acc, const uint32_t* __restrict__ chg_rows, const uint32_t* __restrict__ chg_old,
                      const uint32_t* __restrict__ dctl, const uint32_t* __restrict__ flags, const uint32_t* __restrict__ stop,
                      const AccBatch& bt, uint32_t rows_per_tex) {
    // lock-step batch: rows are global (texture = row / rows_per_tex), the sort key is texture * K1 + centroid
    const uint32_t K = K1 * bt.n_textures;
    // A chunk of listed rows becomes 2 signed entries per row (+x for the new centroid, -x for the old one); the
    // entries are counting-sorted by centroid in shared memory and every half-warp then walks a contiguous slice
    // of the SORTED entries, 8 independent row loads in flight at a time, emitting one 64-bit RED per column at
    // the end of each run of equal centroids.  Chunk size adapts to the list length: long lists use big chunks
    // (long runs, few REDs), short lists small ones (all SMs busy, short dependent chains).
    extern __shared__ uint32_t sm_acc[];
    if (bt.n_textures <= 1 && stop && __ldg(stop)) return;
    if (dctl[33] != 0u) return;  // this pass is a full recompute
    const uint32_t n_use = dctl[32];
    uint32_t chunk_rows = n_use / gridDim.x + 1;
    chunk_rows = chunk_rows < 128u ? 128u : (chunk_rows > (uint32_t)kDeltaRows ? (uint32_t)kDeltaRows : chunk_rows);
    uint32_t* hist = sm_acc;                                          // K
    uint32_t* offs = sm_acc + K;                                      // K
    uint32_t* rows_s = sm_acc + 2 * K;                                // kDeltaRows
    uint16_t* sorted = reinterpret_cast<uint16_t*>(sm_acc + 2 * K + kDeltaRows);   // 2 * kDeltaRows entry ids
    uint16_t* cent_s = sorted + 2 * kDeltaRows;                                    // centroid of each sorted entry
    __shared__ uint32_t warp_tot[kDeltaThreads / 32 + 1];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    unsigned long long* sums_all = reinterpret_cast<unsigned long long*>(acc);
    const uint32_t n_chunks = (n_use + chunk_rows - 1) / chunk_rows;
    constexpr int kPer = kDeltaRows / kDeltaThreads;
    for (uint32_t chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
        const uint32_t first = chunk * chunk_rows;
        const uint32_t len = n_use - first < chunk_rows ? n_use - first : chunk_rows;
        for (uint32_t c = tid; c < K; c += kDeltaThreads) hist[c] = 0;
        __syncthreads();
        uint32_t a_new[kPer], a_old[kPer], r_new[kPer], r_old[kPer];
#pragma unroll
        for (int k = 0; k < kPer; ++k) {
            const uint32_t i = (uint32_t)k * kDeltaThreads + tid;
            a_new[k] = a_old[k] = 0xffffffffu;
            r_new[k] = r_old[k] = 0u;
            if (i < len) {
                const uint32_t row = chg_rows[first + i];
                rows_s[i] = row;
                const uint32_t kbase = (row / rows_per_tex) * K1;
                const uint32_t co = chg_old[first + i], cn = assign[row];
                if (co < K1 && cn < K1 && kbase < K) {
                    a_old[k] = kbase + co;
                    a_new[k] = kbase + cn;
                    r_new[k] = atomicAdd(&hist[a_new[k]], 1u);
                    r_old[k] = atomicAdd(&hist[a_old[k]], 1u);
                }
            }
        }
        __syncthreads();
        {   // exclusive scan of hist -> offs
            const uint32_t per = (K + kDeltaThreads - 1) / kDeltaThreads;
            const uint32_t b0 = tid * per;
            uint32_t local = 0;
            for (uint32_t c = b0; c < b0 + per && c < K; ++c) local += hist[c];
            uint32_t incl = local;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            if (lane == 31) warp_tot[wid] = incl;
            __syncthreads();
            if (wid == 0) {
                uint32_t w = lane < kDeltaThreads / 32 ? warp_tot[lane] : 0u, wi = w;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
                    if (lane >= o) wi += t;
                }
                if (lane < kDeltaThreads / 32) warp_tot[lane] = wi - w;
            }
            __syncthreads();
            uint32_t run = warp_tot[wid] + incl - local;
            for (uint32_t c = b0; c < b0 + per && c < K; ++c) {
                offs[c] = run;
                run += hist[c];
            }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kPer; ++k) {
            const uint32_t i = (uint32_t)k * kDeltaThreads + tid;
            if (a_new[k] < K) {
                const uint32_t p = offs[a_new[k]] + r_new[k];
                sorted[p] = (uint16_t)(2u * i);        // + x
                cent_s[p] = (uint16_t)a_new[k];
            }
            if (a_old[k] < K) {
                const uint32_t p = offs[a_old[k]] + r_old[k];
                sorted[p] = (uint16_t)(2u * i + 1u);   // - x
                cent_s[p] = (uint16_t)a_old[k];
            }
        }
        __syncthreads();
        // half-warp hwid walks sorted entries [p0, p1); lane l of the half-warp owns columns 4l .. 4l+3
        const uint32_t total = offs[K - 1] + hist[K - 1];   // == 2 * len unless an id was out of range (dropped)
        const uint32_t hwid = (uint32_t)tid >> 4, l16 = (uint32_t)tid & 15u;
        constexpr uint32_t kHalfWarps = kDeltaThreads / 16;
        const uint32_t per_hw = (total + kHalfWarps - 1) / kHalfWarps;
        const uint32_t p0 = hwid * per_hw, p1 = p0 + per_hw < total ? p0 + per_hw : total;
        const float4* rows = reinterpret_cast<const float4*>(feat) + l16;
        long long s0 = 0, s1 = 0, s2 = 0, s3 = 0, cnt = 0;
        uint32_t cur = 0xffffffffu;
        float scale = scale1;
        auto flush = [&]() {
            if (cur != 0xffffffffu) {
                const uint32_t tex = cur / K1, cc = cur - tex * K1;
                unsigned long long* sums = sums_all + (size_t)tex * bt.acc_stride;
                unsigned long long* counts = sums + (size_t)K1 * 64;
                unsigned long long* dst = sums + (size_t)cc * 64 + 4 * l16;
                atomicAdd(dst + 0, (unsigned long long)s0);
                atomicAdd(dst + 1, (unsigned long long)s1);
                atomicAdd(dst + 2, (unsigned long long)s2);
                atomicAdd(dst + 3, (unsigned long long)s3);
                if (l16 == 0 && cnt != 0) atomicAdd(counts + cc, (unsigned long long)cnt);
            }
            s0 = s1 = s2 = s3 = cnt = 0;
        };
        for (uint32_t p = p0; p < p1; p += 8) {
            float4 v[8];
            uint32_t e[8], c[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const bool ok = p + q < p1;
                e[q] = ok ? sorted[p + q] : 0u;
                c[q] = ok ? cent_s[p + q] : 0xffffffffu;
                v[q] = ok ? __ldg(rows + (size_t)rows_s[e[q] >> 1] * 16) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                if (p + q < p1) {
                    if (c[q] != cur) {
                        flush();
                        cur = c[q];
                        if (bt.scales) scale = __ldg(bt.scales + cur / K1);
                    }
                    const bool neg = (e[q] & 1u) != 0u;
                    const long long f0 = __float2ll_rn(__fmul_rn(v[q].x, scale)), f1 = __float2ll_rn(__fmul_rn(v[q].y, scale));
                    const long long f2 = __float2ll_rn(__fmul_rn(v[q].z, scale)), f3 = __float2ll_rn(__fmul_rn(v[q].w, scale));
                    s0 += neg ? -f0 : f0;
                    s1 += neg ? -f1 : f1;
                    s2 += neg ? -f2 : f2;
                    s3 += neg ? -f3 : f3;
                    cnt += neg ? -1 : 1;
                }
            }
        }
        flush();
        __syncthreads();
    }
    if (blockIdx.x == 0 && flags) {
        for (uint32_t tex = tid; tex < bt.n_textures; tex += kDeltaThreads) {
            if (bt.n_textures > 1 && stop && __ldg(stop + (size_t)tex * bt.stop_stride)) continue;
            atomicAdd(sums_all + (size_t)tex * bt.acc_stride + (size_t)K1 * 65, (unsigned long long)flags[(size_t)tex * bt.flags_stride]);
        }
    }
}

__global__ void __launch_bounds__(kDeltaThreads)
accumulate_delta_kernel(const float* __restrict__ feat, const uint32_t* __restrict__ assign, uint32_t K1, float scale1,
                        long long* __restrict__ acc, const uint32_t* __restrict__ chg_rows, const uint32_t* __restrict__ chg_old,
                        const uint32_t* __restrict__ dctl, const uint32_t* __restrict__ flags, const uint32_t* __restrict__ stop,
                        AccBatch bt, uint32_t rows_per_tex) {
    accumulate_delta_body(feat, assign, K1, scale1, acc, chg_rows, chg_old, dctl, flags, stop, bt, rows_per_tex);
}

// One launch for the update of a fused Lloyd pass: the mode word dctl[33] (set by the re-rank's last CTA, AccPrep) selects the
// full sort-and-own recompute or the delta update; both bodies return at once in the other mode.  n = rows per texture.
static_assert(kAccThreads == kDeltaThreads, "the combined update kernel runs both bodies with one block shape");
__global__ void __launch_bounds__(kAccThreads, 2)
accumulate_auto_kernel(const float* __restrict__ feat, size_t n, const uint32_t* __restrict__ assign, uint32_t K, float scale1,
                       long long* __restrict__ acc, const uint32_t* __restrict__ chg_rows, const uint32_t* __restrict__ chg_old,
                       const uint32_t* __restrict__ dctl, const uint32_t* __restrict__ flags, const uint32_t* __restrict__ stop,
                       AccBatch bt, uint32_t rows_per_tex) {
    if (dctl[33] != 0u) accumulate_sorted_body(feat, n, assign, K, scale1, acc, flags, stop, bt, dctl);
    else accumulate_delta_body(feat, assign, K, scale1, acc, chg_rows, chg_old, dctl, flags, stop, bt, rows_per_tex);
}

// blockIdx.y = texture of a lock-step batch (strides 0 and inv_scales null for a single texture)
__global__ void finalize_kernel(const long long* __restrict__ acc, uint32_t K, double inv_scale,
                                float* __restrict__ cent, uint32_t* __restrict__ flags, const uint32_t* __restrict__ stop,
                                size_t acc_stride, const double* __restrict__ inv_scales, uint32_t flags_stride,
                                uint32_t stop_stride) {
    if (stop && __ldg(stop + (size_t)blockIdx.y * stop_stride)) return;
    acc += (size_t)blockIdx.y * acc_stride;
    cent += (size_t)blockIdx.y * K * 64;
    flags += (size_t)blockIdx.y * flags_stride;
    if (inv_scales) inv_scale = inv_scales[blockIdx.y];
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)K * 64) return;
    const uint32_t c = (uint32_t)(i >> 6);
    const uint32_t cnt = (uint32_t)acc[(size_t)K * 64 + c];
    if (cnt > 0) {
        const float inv = __fdiv_rn(1.0f, (float)cnt);          // float inv_count = 1.0f / counts[c]
        const float s = (float)((double)acc[i] * inv_scale);    // the sum, rounded to fp32 once
        cent[i] = __fmul_rn(s, inv);                            // newCentroids[c][j] * inv_count
    } else if ((i & 63) == 0) {
        atomicAdd(flags + 1, 1u);
    }
}

// out[0] = number of clusters with count 0 (known before the centroids are overwritten, so the tensor
// engine can compute errors[] only when a reseed will need them)
// see kernels.h: the reference's `if (!hasChanged && iter > N) break;` evaluated on the device
__global__ void lloyd_control_kernel(const long long* __restrict__ acc, uint32_t K, uint32_t iter, uint32_t break_after,
                                     uint32_t* __restrict__ flags, uint32_t* __restrict__ ctl) {
    __shared__ unsigned int total;
    if (ctl[0]) return;
    if (threadIdx.x == 0) total = 0;
    __syncthreads();
    unsigned int local = 0;
    for (uint32_t c = threadIdx.x; c < K; c += blockDim.x) local += acc[(size_t)K * 64 + c] == 0 ? 1u : 0u;
    local = warp_sum(local);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(&total, local);
    __syncthreads();
    if (threadIdx.x == 0) {
        const bool changed = acc[(size_t)K * 65] != 0;
        ctl[1] += 1u;     // assignment passes run
        ctl[2] = total;   // empty clusters of this pass
        if (!changed && iter > break_after) ctl[0] = 1u;  // stop: this pass' update and every later pass are skipped
        flags[0] = 0u;
        flags[1] = 0u;
    }
}

// Lock-step batch: one CTA per texture does what lloyd_control_kernel does for texture blockIdx.x (accumulators,
// flag words and control words are strided per texture).  `sticky` (may be null) is set when a texture that is
// still running has an empty cluster: block-sharded runs cannot reseed on the device (the winner lives on
// another GPU), the host then redoes the batch with the synchronous loop.
__global__ void lloyd_control_batch_kernel(const long long* __restrict__ acc, size_t acc_stride, uint32_t K, uint32_t iter,
                                           uint32_t break_after, uint32_t* __restrict__ flags, uint32_t flags_stride,
                                           uint32_t* __restrict__ ctl_all, uint32_t* __restrict__ sticky) {
    __shared__ unsigned int total;
    acc += (size_t)blockIdx.x * acc_stride;
    flags += (size_t)blockIdx.x * flags_stride;
    uint32_t* ctl = ctl_all + (size_t)blockIdx.x * 3;
    if (ctl[0]) return;
    if (threadIdx.x == 0) total = 0;
    __syncthreads();
    unsigned int local = 0;
    for (uint32_t c = threadIdx.x; c < K; c += blockDim.x) local += acc[(size_t)K * 64 + c] == 0 ? 1u : 0u;
    local = warp_sum(local);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(&total, local);
    __syncthreads();
    if (threadIdx.x == 0) {
        const bool changed = acc[(size_t)K * 65] != 0;
        ctl[1] += 1u;
        ctl[2] = total;
        if (!changed && iter > break_after) ctl[0] = 1u;
        else if (total && sticky) *sticky = 1u;
        flags[0] = 0u;
        flags[1] = 0u;
    }
}

__global__ void count_empty_kernel(const long long* __restrict__ acc, uint32_t K, uint32_t* __restrict__ out) {
    __shared__ unsigned int total;
    if (threadIdx.x == 0) total = 0;
    __syncthreads();
    unsigned int local = 0;
    for (uint32_t c = threadIdx.x; c < K; c += blockDim.x) local += acc[(size_t)K * 64 + c] == 0 ? 1u : 0u;
    local = warp_sum(local);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(&total, local);
    __syncthreads();
    if (threadIdx.x == 0) *out = total;
}

// first index of the maximum of errors[0, n) -- std::max_element semantics -- computed by one CTA.
__device__ unsigned long long cta_argmax_first(const float* errors, size_t n, float* out_val, float* s_val,
                                               unsigned long long* s_idx) {
    float bv = -FLT_MAX;
    unsigned long long bi = ~0ull;
    for (size_t j = threadIdx.x; j < n; j += blockDim.x) {
        const float e = __ldcg(errors + j);
        if (bi == ~0ull || e > bv) {  // ascending strided scan: strict '>' keeps the first maximum
            bv = e;
            bi = j;
        }
    }
    auto merge = [&](float ov, unsigned long long oi) {
        if (oi != ~0ull && (bi == ~0ull || ov > bv || (ov == bv && oi < bi))) {
            bv = ov;
            bi = oi;
        }
    };
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const unsigned long long oi = __shfl_xor_sync(0xffffffffu, bi, o);
        merge(ov, oi);
    }
    __syncthreads();  // protect s_val/s_idx reuse across calls
    if ((threadIdx.x & 31) == 0) {
        s_val[threadIdx.x >> 5] = bv;
        s_idx[threadIdx.x >> 5] = bi;
    }
    __syncthreads();
    if (threadIdx.x < 32) {
        const bool has = threadIdx.x < (blockDim.x >> 5);
        bv = has ? s_val[threadIdx.x] : -FLT_MAX;
        bi = has ? s_idx[threadIdx.x] : ~0ull;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const unsigned long long oi = __shfl_xor_sync(0xffffffffu, bi, o);
            merge(ov, oi);
        }
        if (threadIdx.x == 0) {
            s_idx[32] = bi;
            s_val[32] = bv;
        }
    }
    __syncthreads();
    *out_val = s_val[32];
    return s_idx[32];
}

// Serial semantics of the reference's reseed loop: for ascending c with count 0, take
// the FIRST block with the maximum error, copy its feature vector into the centroid, zero that error.
__global__ void __launch_bounds__(1024) reseed_kernel(const float* __restrict__ feat, size_t n, float* errors,
                                                      const long long* __restrict__ acc, uint32_t K, float* cent,
                                                      uint32_t* flags, const uint32_t* __restrict__ stop) {
    __shared__ float s_val[33];
    __shared__ unsigned long long s_idx[33];
    if (stop && __ldg(stop)) return;
    if (flags[1] == 0) return;  // no empty cluster this pass
    for (uint32_t c = 0; c < K; ++c) {
        if (acc[(size_t)K * 64 + c] != 0) continue;
        float v;
        const unsigned long long worst = cta_argmax_first(errors, n, &v, s_val, s_idx);
        if (threadIdx.x < 64) cent[(size_t)c * 64 + threadIdx.x] = feat[worst * 64 + threadIdx.x];
        if (threadIdx.x == 0) {
            errors[worst] = 0.0f;
            atomicAdd(flags + 2, 1u);
        }
        __threadfence();
        __syncthreads();
    }
}

__global__ void __launch_bounds__(1024) argmax_kernel(const float* __restrict__ errors, size_t n, uint32_t* out2) {
    __shared__ float s_val[33];
    __shared__ unsigned long long s_idx[33];
    float v;
    const unsigned long long worst = cta_argmax_first(errors, n, &v, s_val, s_idx);
    if (threadIdx.x == 0) {
        out2[0] = __float_as_uint(v);
        out2[1] = (uint32_t)worst;
    }
}

__global__ void max_abs_kernel(const float* __restrict__ x, size_t n, uint32_t* out_bits) {
    float m = 0.0f;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float v = fabsf(x[i]);
        if (v != v) m = INFINITY;
        else if (v > m) m = v;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) atomicMax(out_bits, __float_as_uint(m));
}

// accumulate_sorted_kernel's dynamic shared memory can exceed the 48 KB default: opt in once per device to the largest
// size it can ask for (K = kAccMaxK)
cudaError_t accumulate_sorted_optin() {
    static std::atomic<uint64_t> done{0};
    return once_per_device(done, [] {
        return cudaFuncSetAttribute(accumulate_sorted_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)((size_t)kAccMaxK * 8 + (size_t)kAccChunk * 2));
    });
}

}  // namespace

cudaError_t launch_accumulate(const float* feat, size_t n, const uint32_t* assign, uint32_t K, int shift,
                              long long* acc, const uint32_t* flags, const uint32_t* stop, cudaStream_t s, const uint32_t* dctl) {
    if (K <= (uint32_t)kAccMaxK && n >= (size_t)kAccChunk) {
        const size_t smem = (size_t)K * 8 + (size_t)kAccChunk * 2;
        if (cudaError_t e = accumulate_sorted_optin(); e != cudaSuccess) return e;
        size_t chunks = (n + kAccChunk - 1) / kAccChunk;
        size_t grid = chunks < (size_t)YUPVQ_NUM_SMS * 2 ? chunks : (size_t)YUPVQ_NUM_SMS * 2;
        accumulate_sorted_kernel<<<(unsigned)grid, kAccThreads, smem, s>>>(feat, n, assign, K, ldexpf(1.0f, shift), acc, flags, stop,
                                                                          AccBatch{1u, 0, nullptr, 0u, 0u}, dctl);
        return cudaGetLastError();
    }
    size_t blocks = (n * 32 + 255) / 256;
    if (blocks > (size_t)YUPVQ_NUM_SMS * 16) blocks = (size_t)YUPVQ_NUM_SMS * 16;
    if (blocks < 1) blocks = 1;
    accumulate_kernel<<<(unsigned)blocks, 256, 0, s>>>(feat, n, assign, K, ldexpf(1.0f, shift), acc, flags, stop, dctl);
    return cudaGetLastError();
}

cudaError_t launch_finalize_centroids(const long long* acc, uint32_t K, int shift, float* cent, uint32_t* flags,
                                      const uint32_t* stop, cudaStream_t s) {
    finalize_kernel<<<(K * 64 + 255) / 256, 256, 0, s>>>(acc, K, ldexp(1.0, -shift), cent, flags, stop, 0, nullptr, 0u, 0u);
    return cudaGetLastError();
}

cudaError_t launch_acc_prepare(long long* acc, uint32_t K, size_t n, uint32_t force_full, uint32_t* dctl, const uint32_t* stop,
                               cudaStream_t s) {
    if (K > (uint32_t)kDeltaMaxK) force_full = 1u;
    acc_prepare_kernel<<<1, 1024, 0, s>>>(acc, K, (uint32_t)(n / 4), force_full, dctl, stop, 1u, 0);
    return cudaGetLastError();
}

bool delta_batch_supported(uint32_t K, uint32_t n_textures) { return (size_t)K * n_textures <= (size_t)kDeltaMaxK; }

cudaError_t launch_acc_prepare_batch(long long* acc, uint32_t K, size_t total_rows, uint32_t force_full, uint32_t* dctl,
                                     const AccBatch& bt, cudaStream_t s) {
    if (!delta_batch_supported(K, bt.n_textures)) force_full = 1u;
    acc_prepare_kernel<<<1, 1024, 0, s>>>(acc, K, (uint32_t)(total_rows / 4), force_full, dctl, nullptr, bt.n_textures, bt.acc_stride);
    return cudaGetLastError();
}

cudaError_t launch_accumulate_delta(const float* feat, const uint32_t* assign, uint32_t K, int shift, long long* acc,
                                    const uint32_t* chg_rows, const uint32_t* chg_old, const uint32_t* dctl,
                                    const uint32_t* flags, const uint32_t* stop, cudaStream_t s, const AccBatch* batch,
                                    uint32_t rows_per_tex) {
    const AccBatch bt = batch ? *batch : AccBatch{1u, 0, nullptr, 0u, 0u};
    if ((size_t)K * bt.n_textures > (size_t)kDeltaMaxK) return cudaSuccess;  // launch_acc_prepare forces full passes
    const size_t smem = (size_t)K * bt.n_textures * 8 + (size_t)kDeltaRows * 4 + (size_t)kDeltaRows * 4 + (size_t)kDeltaRows * 4;
    static std::atomic<uint64_t> attr_delta{0};  // per device; opt in to the largest size the kernel can ask for
    if (cudaError_t e = once_per_device(attr_delta, [] {
            return cudaFuncSetAttribute(accumulate_delta_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)((size_t)kDeltaMaxK * 8 + (size_t)kDeltaRows * 12));
        }); e != cudaSuccess)
        return e;
    accumulate_delta_kernel<<<YUPVQ_NUM_SMS, kDeltaThreads, smem, s>>>(feat, assign, K, ldexpf(1.0f, shift), acc, chg_rows, chg_old,
                                                                      dctl, flags, stop, bt, batch ? rows_per_tex : 0xffffffffu);
    return cudaGetLastError();
}

// full or delta update in one launch (see accumulate_auto_kernel); batch == null: one texture of n rows
bool accumulate_auto_supported(size_t rows_per_tex, uint32_t K, uint32_t n_textures) {
    return K <= (uint32_t)kAccMaxK && rows_per_tex >= (size_t)kAccChunk && (size_t)K * n_textures <= (size_t)kDeltaMaxK;
}

cudaError_t launch_accumulate_auto(const float* feat, size_t rows_per_tex, const uint32_t* assign, uint32_t K, int shift, long long* acc,
                                   const uint32_t* chg_rows, const uint32_t* chg_old, const uint32_t* dctl, const uint32_t* flags,
                                   const uint32_t* stop, cudaStream_t s, const AccBatch* batch) {
    const AccBatch bt = batch ? *batch : AccBatch{1u, 0, nullptr, 0u, 0u};
    if (!accumulate_auto_supported(rows_per_tex, K, bt.n_textures)) return cudaErrorInvalidValue;
    const size_t smem_full = (size_t)K * 8 + (size_t)kAccChunk * 2;
    const size_t smem_delta = (size_t)K * bt.n_textures * 8 + (size_t)kDeltaRows * 12;
    const size_t smem = smem_full > smem_delta ? smem_full : smem_delta;
    static std::atomic<uint64_t> attr{0};
    if (cudaError_t e = once_per_device(attr, [] {
            return cudaFuncSetAttribute(accumulate_auto_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)((size_t)kDeltaMaxK * 8 + (size_t)kDeltaRows * 12));
        }); e != cudaSuccess)
        return e;
    accumulate_auto_kernel<<<YUPVQ_NUM_SMS * 2, kAccThreads, smem, s>>>(feat, rows_per_tex, assign, K, batch ? 1.0f : ldexpf(1.0f, shift), acc,
                                                                       chg_rows, chg_old, dctl, flags, stop, bt,
                                                                       batch ? (uint32_t)rows_per_tex : 0xffffffffu);
    return cudaGetLastError();
}

bool accumulate_batch_supported(size_t rows_per_tex, uint32_t K) { return K <= (uint32_t)kAccMaxK && rows_per_tex >= (size_t)kAccChunk; }

cudaError_t launch_accumulate_batch(const float* feat, size_t rows_per_tex, const uint32_t* assign, uint32_t K, long long* acc,
                                    const uint32_t* flags, const uint32_t* stop, const AccBatch& bt, cudaStream_t s,
                                    const uint32_t* dctl) {
    if (!accumulate_batch_supported(rows_per_tex, K)) return cudaErrorInvalidValue;
    const size_t smem = (size_t)K * 8 + (size_t)kAccChunk * 2;
    if (cudaError_t e = accumulate_sorted_optin(); e != cudaSuccess) return e;
    size_t chunks = (rows_per_tex + kAccChunk - 1) / kAccChunk * bt.n_textures;
    size_t grid = chunks < (size_t)YUPVQ_NUM_SMS * 2 ? chunks : (size_t)YUPVQ_NUM_SMS * 2;
    accumulate_sorted_kernel<<<(unsigned)grid, kAccThreads, smem, s>>>(feat, rows_per_tex, assign, K, 1.0f, acc, flags, stop, bt, dctl);
    return cudaGetLastError();
}

cudaError_t launch_finalize_centroids_batch(const long long* acc, size_t acc_stride, uint32_t n_textures, uint32_t K,
                                            const double* inv_scales, float* cent, uint32_t* flags, uint32_t flags_stride,
                                            const uint32_t* stop, uint32_t stop_stride, cudaStream_t s) {
    finalize_kernel<<<dim3((K * 64 + 255) / 256, n_textures), 256, 0, s>>>(acc, K, 1.0, cent, flags, stop, acc_stride, inv_scales,
                                                                           flags_stride, stop_stride);
    return cudaGetLastError();
}

cudaError_t launch_count_empty(const long long* acc, uint32_t K, uint32_t* out, cudaStream_t s) {
    count_empty_kernel<<<1, 256, 0, s>>>(acc, K, out);
    return cudaGetLastError();
}

cudaError_t launch_lloyd_control(const long long* acc, uint32_t K, uint32_t iter, uint32_t break_after, uint32_t* flags,
                                 uint32_t* ctl, cudaStream_t s) {
    lloyd_control_kernel<<<1, 256, 0, s>>>(acc, K, iter, break_after, flags, ctl);
    return cudaGetLastError();
}

cudaError_t launch_lloyd_control_batch(const long long* acc, size_t acc_stride, uint32_t n_textures, uint32_t K, uint32_t iter,
                                       uint32_t break_after, uint32_t* flags, uint32_t flags_stride, uint32_t* ctl,
                                       uint32_t* sticky, cudaStream_t s) {
    lloyd_control_batch_kernel<<<n_textures, 256, 0, s>>>(acc, acc_stride, K, iter, break_after, flags, flags_stride, ctl, sticky);
    return cudaGetLastError();
}

cudaError_t launch_reseed_empty(const float* feat, size_t n, float* errors, const long long* acc, uint32_t K,
                                float* cent, uint32_t* flags, const uint32_t* stop, cudaStream_t s) {
    reseed_kernel<<<1, 1024, 0, s>>>(feat, n, errors, acc, K, cent, flags, stop);
    return cudaGetLastError();
}

cudaError_t launch_argmax_error(const float* errors, size_t n, uint32_t* out2, cudaStream_t s) {
    argmax_kernel<<<1, 1024, 0, s>>>(errors, n, out2);
    return cudaGetLastError();
}

cudaError_t launch_max_abs(const float* x, size_t n, uint32_t* out_bits, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    size_t blocks = (n + 255) / 256;
    if (blocks > (size_t)YUPVQ_NUM_SMS * 8) blocks = (size_t)YUPVQ_NUM_SMS * 8;
    max_abs_kernel<<<(unsigned)blocks, 256, 0, s>>>(x, n, out_bits);
    return cudaGetLastError();
}

}  // namespace yupvq
