// A21: the SAD-metric Lloyd path (vq_Metric == SAD), reference src/vq_encoder.cpp:380-430 (BuildCodebook) and
// :473-484 (QuantizeBlocks).  Everything on this path is integer arithmetic, so the GPU result is bit-identical to
// the reference for the same seeds: distance = sum |a - b| over the 16*ch pixel bytes (BlockDistanceSAD, :201-208),
// argmin with strict '<' in ascending centroid order, centroid byte = uint64 sum / count (truncating, :421).
//
// Data layout: the blocks are the uint4 pixel planes the K-Means++ seeding already reads (plane q = bytes
// [16q, 16q + 16) of every block); the centroids are K pixel blocks of 16*ch bytes, i.e. ch uint4 per centroid.
//   sad_assign_kernel   one thread = TWO blocks held in registers (2 * ch uint4); the centroid table is staged in
//                       shared memory in chunks of 512 and read as 16-byte broadcasts, each word feeds two
//                       VABSDIFF4.U8.ACC (the native 4-byte SAD-accumulate of sm_100a).  Integer-ALU bound:
//                       N * K * 4 * ch VABSDIFF4 per pass; HBM traffic is 16*ch + 8 bytes per block.
//   update              the pixel bytes are expanded once per call to an N x 64 fp32 matrix (integer-valued), which
//                       lets the pass reuse the sort-and-own accumulation kernel (update.cu) with shift 0: int64
//                       sums of integers are exact, sad_finalize_kernel does the truncating division.
#include <cfloat>

#include "common.cuh"
#include "kernels.h"

namespace yupvq {

namespace {

constexpr int kSadThreads = 256;
constexpr int kSadChunk = 512;  // centroids staged in shared memory at a time

// VABSDIFF4.U8.ACC with the accumulate operand spelled out: `__vsadu4(a, b) + acc` compiles to a zero-accumulate
// VABSDIFF4 plus an IADD3 tree, i.e. ~1.6 integer instructions per four byte pairs instead of one.
__device__ __forceinline__ uint32_t sad4_acc(uint32_t a, uint32_t b, uint32_t acc) {
    uint32_t d;
    asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(acc));
    return d;
}

__device__ __forceinline__ uint32_t sad16(const uint4& a, const uint4& b, uint32_t acc) {
    acc = sad4_acc(a.x, b.x, acc);
    acc = sad4_acc(a.y, b.y, acc);
    acc = sad4_acc(a.z, b.z, acc);
    acc = sad4_acc(a.w, b.w, acc);
    return acc;
}

template <int CH>
__global__ void __launch_bounds__(kSadThreads) sad_assign_kernel(const uint4* __restrict__ planes, size_t stride, size_t n,
                                                                const uint4* __restrict__ cent, uint32_t K,
                                                                uint32_t* __restrict__ assign, float* __restrict__ errors,
                                                                uint32_t* __restrict__ flags, const uint32_t* __restrict__ stop) {
    __shared__ uint4 s_cent[kSadChunk * CH];
    if (stop && __ldg(stop)) return;
    const size_t pairs = (n + 1) / 2;
    const size_t per_cta = (size_t)kSadThreads;
    for (size_t p0 = (size_t)blockIdx.x * per_cta; p0 < pairs; p0 += (size_t)gridDim.x * per_cta) {
        // blocks b0 = p, b1 = p + pairs: both halves of the block range are read coalesced
        const size_t p = p0 + threadIdx.x;
        const bool v0 = p < pairs, v1 = p + pairs < n;
        uint4 x0[CH], x1[CH];
#pragma unroll
        for (int q = 0; q < CH; ++q) {
            x0[q] = v0 ? ldg_nc_v4(planes + q * stride + p) : make_uint4(0, 0, 0, 0);
            x1[q] = v1 ? ldg_nc_v4(planes + q * stride + p + pairs) : make_uint4(0, 0, 0, 0);
        }
        uint32_t best0 = 0xffffffffu, best1 = 0xffffffffu, c0 = 0, c1 = 0;
        for (uint32_t base = 0; base < K; base += kSadChunk) {
            const uint32_t m = K - base < (uint32_t)kSadChunk ? K - base : (uint32_t)kSadChunk;
            __syncthreads();
            for (uint32_t i = threadIdx.x; i < m * CH; i += kSadThreads) s_cent[i] = __ldg(cent + (size_t)base * CH + i);
            __syncthreads();
#pragma unroll 2
            for (uint32_t c = 0; c < m; ++c) {
                uint32_t d0 = 0, d1 = 0;
#pragma unroll
                for (int q = 0; q < CH; ++q) {
                    const uint4 cv = s_cent[c * CH + q];
                    d0 = sad16(x0[q], cv, d0);
                    d1 = sad16(x1[q], cv, d1);
                }
                // strict '<' in ascending order keeps the first minimum (:389, :480); float(d) is exact (d <= 16320)
                if (d0 < best0) { best0 = d0; c0 = base + c; }
                if (d1 < best1) { best1 = d1; c1 = base + c; }
            }
        }
        bool changed = false;
        if (v0) {
            changed |= assign[p] != c0;
            assign[p] = c0;
            if (errors) errors[p] = (float)best0;
        }
        if (v1) {
            changed |= assign[p + pairs] != c1;
            assign[p + pairs] = c1;
            if (errors) errors[p + pairs] = (float)best1;
        }
        if (flags && __any_sync(0xffffffffu, changed) && (threadIdx.x & 31) == 0) flags[0] = 1u;
    }
}

// pixel planes -> N x 64 fp32 rows (byte j of the block at column j, columns >= 16*ch zero)
template <int CH>
__global__ void __launch_bounds__(256) sad_expand_kernel(const uint4* __restrict__ planes, size_t stride, size_t n,
                                                        float* __restrict__ rows) {
    // 16 threads per block: thread j writes columns [4j, 4j + 4) = one 32-bit word of plane j / 4
    const size_t i = (size_t)blockIdx.x * 16 + (threadIdx.x >> 4);
    const int j = threadIdx.x & 15;
    if (i >= n) return;
    float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
    if (j < 4 * CH) {
        const uint32_t w = reinterpret_cast<const uint32_t*>(planes + (j >> 2) * stride + i)[j & 3];
        o = make_float4((float)(w & 0xffu), (float)((w >> 8) & 0xffu), (float)((w >> 16) & 0xffu), (float)(w >> 24));
    }
    reinterpret_cast<float4*>(rows)[i * 16 + j] = o;
}

// K x 64 fp32 (integer-valued) -> K pixel blocks of 16*ch bytes
__global__ void sad_pack_kernel(const float* __restrict__ cent_f, uint32_t K, int ch, uint8_t* __restrict__ cent_px) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= K * 16u * (uint32_t)ch) return;
    const uint32_t c = i / (16u * ch), j = i % (16u * ch);
    cent_px[i] = (uint8_t)cent_f[(size_t)c * 64 + j];
}

// centroid byte = sum / count, truncating (:421); clusters with count 0 keep their centroid and are counted in
// flags[1] for the reseed
__global__ void sad_finalize_kernel(const long long* __restrict__ acc, uint32_t K, float* __restrict__ cent_f,
                                    uint32_t* __restrict__ flags, const uint32_t* __restrict__ stop) {
    if (stop && __ldg(stop)) return;
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)K * 64) return;
    const uint32_t c = (uint32_t)(i >> 6);
    const unsigned long long cnt = (unsigned long long)acc[(size_t)K * 64 + c];
    if (cnt > 0) cent_f[i] = (float)(uint32_t)((unsigned long long)acc[i] / cnt);
    else if ((i & 63) == 0) atomicAdd(flags + 1, 1u);
}

}  // namespace

cudaError_t launch_sad_assign(const uint4* planes, size_t stride, size_t n, int ch, const uint8_t* cent_px, uint32_t K,
                              uint32_t* assign, float* errors, uint32_t* flags, const uint32_t* stop, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    size_t grid = ((n + 1) / 2 + kSadThreads - 1) / kSadThreads;
    if (grid > (size_t)YUPVQ_NUM_SMS * 8) grid = (size_t)YUPVQ_NUM_SMS * 8;
    const uint4* c4 = reinterpret_cast<const uint4*>(cent_px);
    switch (ch) {
        case 1: sad_assign_kernel<1><<<(unsigned)grid, kSadThreads, 0, s>>>(planes, stride, n, c4, K, assign, errors, flags, stop); break;
        case 2: sad_assign_kernel<2><<<(unsigned)grid, kSadThreads, 0, s>>>(planes, stride, n, c4, K, assign, errors, flags, stop); break;
        case 3: sad_assign_kernel<3><<<(unsigned)grid, kSadThreads, 0, s>>>(planes, stride, n, c4, K, assign, errors, flags, stop); break;
        case 4: sad_assign_kernel<4><<<(unsigned)grid, kSadThreads, 0, s>>>(planes, stride, n, c4, K, assign, errors, flags, stop); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

cudaError_t launch_sad_expand(const uint4* planes, size_t stride, size_t n, int ch, float* rows, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    const unsigned grid = (unsigned)((n + 15) / 16);
    switch (ch) {
        case 1: sad_expand_kernel<1><<<grid, 256, 0, s>>>(planes, stride, n, rows); break;
        case 2: sad_expand_kernel<2><<<grid, 256, 0, s>>>(planes, stride, n, rows); break;
        case 3: sad_expand_kernel<3><<<grid, 256, 0, s>>>(planes, stride, n, rows); break;
        case 4: sad_expand_kernel<4><<<grid, 256, 0, s>>>(planes, stride, n, rows); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

cudaError_t launch_sad_pack(const float* cent_f, uint32_t K, int ch, uint8_t* cent_px, cudaStream_t s) {
    const uint32_t total = K * 16u * (uint32_t)ch;
    sad_pack_kernel<<<(total + 255) / 256, 256, 0, s>>>(cent_f, K, ch, cent_px);
    return cudaGetLastError();
}

cudaError_t launch_sad_finalize(const long long* acc, uint32_t K, float* cent_f, uint32_t* flags, const uint32_t* stop,
                                cudaStream_t s) {
    sad_finalize_kernel<<<(K * 64 + 255) / 256, 256, 0, s>>>(acc, K, cent_f, flags, stop);
    return cudaGetLastError();
}

}  // namespace yupvq
