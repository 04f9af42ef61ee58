// Shared helpers for the sm_100a VQ kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define YUPVQ_D 64            // floats per block feature vector: 16 px x [L, a, b, A] (reference include/vq_encoder.h:16-20)
#define YUPVQ_NUM_SMS 148     // B200

// BCFormat values of the reference's include/vq_bcn_types.h:27-36
enum { YUPVQ_FMT_BC1 = 1, YUPVQ_FMT_BC2 = 2, YUPVQ_FMT_BC3 = 3, YUPVQ_FMT_BC4 = 4, YUPVQ_FMT_BC5 = 5, YUPVQ_FMT_BC6H = 6, YUPVQ_FMT_BC7 = 7 };

namespace yupvq {

__device__ __forceinline__ uint4 ldg_nc_v4(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

__device__ __forceinline__ float4 ldg_nc_f4(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Exact restatement of the reference's weighted squared distance between two 64-float Oklab
// blocks, reference src/vq_encoder.cpp:212-237: eight lane accumulators,
// sub -> mul by weight (2 for L, exact) -> fma, then (lane j + lane j+4) and two hadd levels.
// Round-to-nearest intrinsics keep nvcc from contracting or reassociating anything.
struct DistAcc {
    float a[8];
    __device__ __forceinline__ void reset() {
#pragma unroll
        for (int l = 0; l < 8; ++l) a[l] = 0.0f;
    }
    // i = 0..7 selects floats [8i, 8i+8): x and c are the two pixels' [L a b A L a b A]
    __device__ __forceinline__ void step(const float4& x0, const float4& x1, const float4& c0, const float4& c1) {
        float d;
        d = __fsub_rn(x0.x, c0.x); d = __fmul_rn(d, 2.0f); a[0] = __fmaf_rn(d, d, a[0]);
        d = __fsub_rn(x0.y, c0.y);                         a[1] = __fmaf_rn(d, d, a[1]);
        d = __fsub_rn(x0.z, c0.z);                         a[2] = __fmaf_rn(d, d, a[2]);
        d = __fsub_rn(x0.w, c0.w);                         a[3] = __fmaf_rn(d, d, a[3]);
        d = __fsub_rn(x1.x, c1.x); d = __fmul_rn(d, 2.0f); a[4] = __fmaf_rn(d, d, a[4]);
        d = __fsub_rn(x1.y, c1.y);                         a[5] = __fmaf_rn(d, d, a[5]);
        d = __fsub_rn(x1.z, c1.z);                         a[6] = __fmaf_rn(d, d, a[6]);
        d = __fsub_rn(x1.w, c1.w);                         a[7] = __fmaf_rn(d, d, a[7]);
    }
    __device__ __forceinline__ float finish() const {
        const float s0 = __fadd_rn(a[0], a[4]), s1 = __fadd_rn(a[1], a[5]);
        const float s2 = __fadd_rn(a[2], a[6]), s3 = __fadd_rn(a[3], a[7]);
        return __fadd_rn(__fadd_rn(s0, s1), __fadd_rn(s2, s3));
    }
};

// distance between a block held in registers (16 float4) and a centroid in shared/global memory
__device__ __forceinline__ float oklab_dist_reg(const float4 (&x)[16], const float4* __restrict__ c) {
    DistAcc acc;
    acc.reset();
#pragma unroll
    for (int i = 0; i < 8; ++i) acc.step(x[2 * i], x[2 * i + 1], c[2 * i], c[2 * i + 1]);
    return acc.finish();
}

}  // namespace yupvq
