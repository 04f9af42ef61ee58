// K3 (exact engine): nearest-centroid assignment in fp32 with the reference's operation order.
// Reference: the loops of src/vq_encoder.cpp:330-340, :462-471, :561-571, :643-656 calling
// OklabBlockDistanceSq_SIMD (:212-237); argmin with strict '<' so the lowest index wins ties.
// This kernel is the correctness anchor: it reproduces indices AND distances bit for bit.  The
// tcgen05 front-end (assign_tc.cu) narrows each block to a short candidate list and reuses the same
// arithmetic (common.cuh DistAcc) for the exact re-rank.
// Roofline: fp32-ALU bound (N*K*64 sub + fma); 256 B read + 8 B written per block.
#include <cfloat>

#include "common.cuh"
#include "kernels.h"

namespace yupvq {

namespace {

constexpr int kTile = 128;   // blocks (== threads) per CTA
constexpr int kPitch4 = 17;  // float4 per shared-memory row
constexpr int kCTile = 32;   // centroids staged per step

__global__ void __launch_bounds__(kTile) assign_exact_kernel(const float* __restrict__ feat, size_t n,
                                                             const float* __restrict__ cent, uint32_t K,
                                                             uint32_t* __restrict__ assign, float* __restrict__ errors,
                                                             uint32_t* __restrict__ flags, const uint32_t* __restrict__ stop) {
    __shared__ float4 xs[kTile * kPitch4];
    if (stop && __ldg(stop)) return;
    __shared__ float4 cs[kCTile * 16];
    const size_t tile_first = (size_t)blockIdx.x * kTile;
    const size_t rows = (n - tile_first) < (size_t)kTile ? (n - tile_first) : (size_t)kTile;
    const float4* src = reinterpret_cast<const float4*>(feat) + tile_first * 16;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int idx = threadIdx.x + kTile * k;
        const int row = idx >> 4, col = idx & 15;
        if ((size_t)row < rows) xs[row * kPitch4 + col] = ldg_nc_f4(src + idx);
    }
    __syncthreads();
    const bool valid = threadIdx.x < rows;
    float4 x[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) x[k] = valid ? xs[threadIdx.x * kPitch4 + k] : make_float4(0.f, 0.f, 0.f, 0.f);

    float best = FLT_MAX;
    uint32_t best_c = 0;
    for (uint32_t c0 = 0; c0 < K; c0 += kCTile) {
        const uint32_t cnt = (K - c0) < (uint32_t)kCTile ? (K - c0) : (uint32_t)kCTile;
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < cnt * 16; i += kTile)
            cs[i] = reinterpret_cast<const float4*>(cent)[(size_t)c0 * 16 + i];
        __syncthreads();
        for (uint32_t cc = 0; cc < cnt; ++cc) {
            const float d = oklab_dist_reg(x, &cs[cc * 16]);
            if (d < best) {
                best = d;
                best_c = c0 + cc;
            }
        }
    }
    if (valid) {
        const size_t blk = tile_first + threadIdx.x;
        if (flags && assign[blk] != best_c) flags[0] = 1u;  // hasChanged (benign race: all writers store 1)
        assign[blk] = best_c;
        if (errors) errors[blk] = best;
    }
}

__global__ void exact_errors_kernel(const float* __restrict__ feat, size_t n, const float* __restrict__ cent,
                                    const uint32_t* __restrict__ assign, float* __restrict__ errors,
                                    const uint32_t* __restrict__ stop, const uint32_t* __restrict__ n_empty) {
    // launched every pass but needed only when a cluster came out empty: a small grid keeps the early exit cheap
    if ((stop && __ldg(stop)) || (n_empty && __ldg(n_empty) == 0)) return;
    for (size_t blk = (size_t)blockIdx.x * blockDim.x + threadIdx.x; blk < n; blk += (size_t)gridDim.x * blockDim.x) {
        const float4* x = reinterpret_cast<const float4*>(feat) + blk * 16;
        const float4* c = reinterpret_cast<const float4*>(cent) + (size_t)assign[blk] * 16;
        DistAcc acc;
        acc.reset();
#pragma unroll
        for (int i = 0; i < 8; ++i) acc.step(x[2 * i], x[2 * i + 1], c[2 * i], c[2 * i + 1]);
        errors[blk] = acc.finish();
    }
}

}  // namespace

cudaError_t launch_assign_exact(const float* feat, size_t n, const float* cent, uint32_t K, uint32_t* assign,
                                float* errors, uint32_t* flags, const uint32_t* stop, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    assign_exact_kernel<<<(unsigned)((n + kTile - 1) / kTile), kTile, 0, s>>>(feat, n, cent, K, assign, errors, flags, stop);
    return cudaGetLastError();
}

cudaError_t launch_exact_errors(const float* feat, size_t n, const float* cent, const uint32_t* assign, float* errors,
                                const uint32_t* stop, const uint32_t* n_empty, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    size_t grid = (n + 127) / 128;
    if (grid > (size_t)YUPVQ_NUM_SMS * 8) grid = (size_t)YUPVQ_NUM_SMS * 8;
    exact_errors_kernel<<<(unsigned)grid, 128, 0, s>>>(feat, n, cent, assign, errors, stop, n_empty);
    return cudaGetLastError();
}

}  // namespace yupvq
