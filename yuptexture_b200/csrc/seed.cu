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

#include "common.cuh"
#include "kernels.h"

namespace yupvq {

namespace {

constexpr int kThreads = 256;
constexpr int kMaxGrid = YUPVQ_NUM_SMS * 8;

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
            sad = __vsadu4(v.x, sp[q].x) + sad;
            sad = __vsadu4(v.y, sp[q].y) + sad;
            sad = __vsadu4(v.z, sp[q].z) + sad;
            sad = __vsadu4(v.w, sp[q].w) + sad;
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

inline size_t chunk_for(size_t n, int grid) {
    size_t c = (n + grid - 1) / grid;
    return (c + kThreads - 1) / kThreads * kThreads;
}

}  // namespace

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

cudaError_t launch_seed_round_hdr(const SeedState& st, const float* feat, size_t n, uint32_t round, uint32_t K,
                                  cudaStream_t s) {
    (void)K;
    const size_t chunk = chunk_for(n, st.grid);
    seed_round_hdr_kernel<<<st.grid, kThreads, 0, s>>>(st, feat, n, chunk, round);
    return cudaGetLastError();
}

}  // namespace yupvq
