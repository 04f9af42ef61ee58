// Operand preparation shared by assign_tc.cu (prep_centroids_kernel) and tail.cu (the fused per-pass tail): one warp
// turns one centroid into its row of the tensor front-end's B operand.  See assign_tc.cu's header for the layout.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace yupvq {

// two-term bf16 split: v = hi + lo + r, |r| <= 2^-18 |v| (v - hi is exact in fp32)
__device__ __forceinline__ void split_bf16(float v, uint16_t& hi, uint16_t& lo) {
    const __nv_bfloat16 h = __float2bfloat16_rn(v);
    const __nv_bfloat16 l = __float2bfloat16_rn(__fsub_rn(v, __bfloat162float(h)));
    hi = __bfloat16_as_ushort(h);
    lo = __bfloat16_as_ushort(l);
}

// B' row (256 B) = [bf16 hi of w (c - mu) x 64 | bf16 lo x 64]; the ext row (32 B = 16 bf16) holds a three-term bf16
// split of -|c'|^2 / 2 (24 bits); rows K..K_pad-1 are padding that can never win (score -1e30).  cmax = max_c |c'|
// (rounded up).  One warp per row: lane owns columns `lane` and `lane + 32`; x0 / x1 are the centroid's values in those
// columns (ignored when !real), mu the texture's mean row.
//   rgb: the 48 colour columns are packed first; the alpha column is dropped from the product (x' is exactly 0 there) but
//   stays in the norm, which is the true |c'|^2; the folded norm terms sit in columns 48..50 of the hi half.
__device__ __forceinline__ void tc_prep_centroid_warp(float x0, float x1, bool real, const float* __restrict__ mu, uint16_t* __restrict__ brow,
                                                      uint16_t* __restrict__ bext_row, unsigned int* __restrict__ cmax_bits, int rgb, int lane) {
    float v[2] = {0.0f, 0.0f};
    if (real) {
        const int c0 = lane, c1 = lane + 32;
        v[0] = ((c0 & 3) == 0 ? 2.0f : 1.0f) * (x0 - mu[c0]);
        v[1] = ((c1 & 3) == 0 ? 2.0f : 1.0f) * (x1 - mu[c1]);
        if (!isfinite(v[0])) v[0] = 0.0f;   // a non-finite centroid is handled by the exact scan (its blocks overflow)
        if (!isfinite(v[1])) v[1] = 0.0f;
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int i = lane + 32 * h;
        uint16_t hi16, lo16;
        split_bf16(v[h], hi16, lo16);
        if (rgb) {
            if ((i & 3) != 3) {
                brow[3 * (i >> 2) + (i & 3)] = hi16;
                brow[64 + 3 * (i >> 2) + (i & 3)] = lo16;
            } else {
                brow[112 + (i >> 2)] = 0;
            }
        } else {
            brow[i] = hi16;
            brow[64 + i] = lo16;
        }
    }
    // |c'|^2 in fp64: columns 0..31 and 32..63 reduced separately (butterfly), then added
    double s0 = (double)v[0] * (double)v[0], s1 = (double)v[1] * (double)v[1];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s0 += __shfl_xor_sync(0xffffffffu, s0, o);
        s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    }
    const double ss = s0 + s1;
    if (lane < 16) {
        float e = 0.0f;
        if (real) {
            const double nh = -0.5 * ss;
            const float hi = __bfloat162float(__float2bfloat16_rn((float)nh));
            const float mid = __bfloat162float(__float2bfloat16_rn((float)(nh - (double)hi)));
            const float lo = __bfloat162float(__float2bfloat16_rn((float)(nh - (double)hi - (double)mid)));
            e = lane == 0 ? hi : (lane == 1 ? mid : (lane == 2 ? lo : 0.0f));
        } else if (lane == 0) {
            e = -1.0e30f;
        }
        const uint16_t eb = __bfloat16_as_ushort(__float2bfloat16_rn(e));
        bext_row[lane] = eb;
        if (rgb) brow[48 + lane] = eb;   // folded norm inside the hi half
    }
    if (lane == 0 && real) {
        const float nrm = (float)sqrt(ss) * 1.0001f;
        atomicMax(cmax_bits, __float_as_uint(nrm));  // non-negative floats order like their bit patterns
    }
}

}  // namespace yupvq
