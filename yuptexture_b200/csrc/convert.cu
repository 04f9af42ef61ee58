// K1 / K1h / K6: BCn blocks, uint8 pixel blocks, float tiles  ->  64-float Oklab feature vectors
// (+ the uint8 pixel planes the LDR K-Means++ seeding reads), and the inverse conversion of the K
// centroids.  Reference: src/vq_bcn_compressor.cpp:142-146 (per-block decode), :235-252 (HDR tile
// gather), src/vq_encoder.cpp:37-194 (PixelToOklab / OklabToPixel / RgbaFloatToOklab /
// OklabToRgbaFloat and their block wrappers).
//
// HBM layout: features are row-major N x 64 fp32 (256 B per block, written through a
// shared-memory transpose so every warp store is a full 512-B run); pixel blocks are `ch` planes of
// uint4.  Roofline: HBM-bound; algorithmic bytes per block = blockSize read + 256 B features
// + 16*ch B pixels written.
#include <cmath>

#include "bcn_decode.cuh"
#include "colour.cuh"
#include "kernels.h"
#include "tc_prep.cuh"

namespace yupvq {

namespace {

constexpr int kTile = 128;       // blocks per CTA == threads per CTA
constexpr int kPitch4 = 17;      // float4 per shared-memory row (16 + 1 pad)

// ---- host: sRGB8 -> linear table with the host libm (the function the reference itself calls) ----
double host_srgb8_to_linear(double value) {  // colorm.h:4023-4041
    if (value > 0.0) {
        value /= 255.0;
        if (value <= 0.04045) return value / 12.92;
        return std::pow((value + 0.055) / 1.055, 2.4);
    }
    value = -value;
    value /= 255.0;
    if (value <= 0.04045) return -(value / 12.92);
    return -(std::pow((value + 0.055) / 1.055, 2.4));
}

// ---- per-block feature construction ------------------------------------------------------------

template <int CH>
__device__ __forceinline__ uint32_t texel_byte(const uint32_t (&w)[4 * CH], int p, int c) {
    const int b = p * CH + c;
    return (w[b >> 2] >> (8 * (b & 3))) & 255u;
}

// VQEncoder::PixelBlockToOklabBlock (src/vq_encoder.cpp:96-102) on a pixel block held as packed
// words.  A BC1/BC3 block has at most four distinct colours, so the conversions are de-duplicated, in three
// steps that keep the warp converged through the fp64 code: (1) integer only: every texel gets the slot of its
// colour in a 4-entry key list, or an overflow bit once the block has shown a fifth colour; (2) slot j is
// converted by all lanes together while any lane of the warp has more than j colours; (3) texels pick their slot,
// overflow texels are converted one by one (BC7 / raw pixel blocks with many colours, no worse than a plain loop).
template <int CH>
__device__ __forceinline__ uint32_t texel_rgb_key(const uint32_t (&w)[4 * CH], int p) {
    if (CH <= 2) return texel_byte<CH>(w, p, 0) * 0x010101u;
    return texel_byte<CH>(w, p, 0) | (texel_byte<CH>(w, p, 1) << 8) | (texel_byte<CH>(w, p, 2) << 16);
}

template <int CH>
__device__ __forceinline__ void block_features(const uint32_t (&w)[4 * CH], float4 (&f)[16]) {
    uint32_t key[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};  // never equals a 24-bit key
    uint32_t slots = 0, overflow = 0;
    int count = 0;
#pragma unroll
    for (int p = 0; p < 16; ++p) {
        const uint32_t k = texel_rgb_key<CH>(w, p);
        uint32_t s;
        if (k == key[0]) s = 0;
        else if (k == key[1]) s = 1;
        else if (k == key[2]) s = 2;
        else if (k == key[3]) s = 3;
        else if (count < 4) {
            s = (uint32_t)count;
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (j == count) key[j] = k;
            ++count;
        } else {
            s = 0;
            overflow |= 1u << p;
        }
        slots |= s << (2 * p);
    }
    float3 val[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        val[j] = make_float3(0.f, 0.f, 0.f);
        if (__any_sync(__activemask(), count > j)) {
            const uint32_t k = count > j ? key[j] : 0u;
            val[j] = srgb8_to_oklab(k & 255u, (k >> 8) & 255u, (k >> 16) & 255u);
        }
    }
#pragma unroll
    for (int p = 0; p < 16; ++p) {
        const uint32_t s = (slots >> (2 * p)) & 3u;
        float3 o = s == 0 ? val[0] : s == 1 ? val[1] : s == 2 ? val[2] : val[3];
        if (overflow & (1u << p)) {
            const uint32_t k = texel_rgb_key<CH>(w, p);
            o = srgb8_to_oklab(k & 255u, (k >> 8) & 255u, (k >> 16) & 255u);
        }
        float alpha = 1.0f;
        if (CH == 2) alpha = __fdiv_rn((float)texel_byte<CH>(w, p, 1), 255.0f);
        if (CH == 4) alpha = __fdiv_rn((float)texel_byte<CH>(w, p, 3), 255.0f);
        f[p] = make_float4(o.x, o.y, o.z, alpha);
    }
}

// The same features for a block whose colours come from a 4-entry palette (BC1, BC3 colour part): the four
// palette colours are converted by every lane at the same point of the program, then selected per texel by the
// 2-bit index.  block_features' cache makes the same four conversions per block, but the lanes of a warp miss at
// different texels, so the warp walks the fp64 conversion at nearly all sixteen positions; here it walks it four
// times.  Same function on the same bytes, so the features are bit-identical.
template <int CH>
__device__ __forceinline__ void palette_features(const uint32_t (&pal)[4], uint32_t sel, const uint32_t (&w)[4 * CH],
                                                 float4 (&f)[16]) {
    float3 val[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        uint32_t r = pal[j] & 255u, g = (pal[j] >> 8) & 255u, b = (pal[j] >> 16) & 255u;
        if (CH == 1) r = g = b;       // 1-channel output of BC1/BC3 takes B (pack_block)
        else if (CH == 2) g = b = r;  // 2-channel output takes (R, G): colour from R, alpha from G
        val[j] = srgb8_to_oklab(r, g, b);
    }
#pragma unroll
    for (int p = 0; p < 16; ++p) {
        const uint32_t k = (sel >> (2 * p)) & 3u;
        const float3 o = k == 0 ? val[0] : k == 1 ? val[1] : k == 2 ? val[2] : val[3];
        float alpha = 1.0f;
        if (CH == 2) alpha = __fdiv_rn((float)texel_byte<CH>(w, p, 1), 255.0f);
        if (CH == 4) alpha = __fdiv_rn((float)texel_byte<CH>(w, p, 3), 255.0f);
        f[p] = make_float4(o.x, o.y, o.z, alpha);
    }
}

// RGBA texels -> packed pixel block with the reference decoder's channel selection (see
// oracle/vq_oracle.cpp store_pixels: 1-channel output takes B (R for BC4), 2-channel takes (R,G)
// except BC5 -> (G,B)).
template <int FMT, int CH>
__device__ __forceinline__ void pack_block(const uint32_t (&t)[16], uint32_t (&w)[4 * CH]) {
#pragma unroll
    for (int i = 0; i < 4 * CH; ++i) w[i] = 0;
#pragma unroll
    for (int p = 0; p < 16; ++p) {
#pragma unroll
        for (int c = 0; c < CH; ++c) {
            int src = c;
            if (CH == 1) src = (FMT == YUPVQ_FMT_BC4) ? 0 : 2;
            if (CH == 2 && FMT == YUPVQ_FMT_BC5) src = c + 1;
            const uint32_t v = (t[p] >> (8 * src)) & 255u;
            const int b = p * CH + c;
            w[b >> 2] |= v << (8 * (b & 3));
        }
    }
}

template <int CH>
__device__ __forceinline__ void store_planes(const uint32_t (&w)[4 * CH], uint4* planes, size_t stride, size_t blk) {
#pragma unroll
    for (int q = 0; q < CH; ++q) planes[q * stride + blk] = make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
}

// shared-memory transpose: thread t holds block row t; the CTA then streams the tile out with
// consecutive threads writing consecutive float4 (512 B per warp store).
__device__ __forceinline__ void store_features_tile(const float4 (&f)[16], bool valid, float4* tile, float* feat,
                                                    size_t tile_first, size_t n) {
    if (valid) {
#pragma unroll
        for (int k = 0; k < 16; ++k) tile[threadIdx.x * kPitch4 + k] = f[k];
    }
    __syncthreads();
    const size_t rows = (n - tile_first) < (size_t)kTile ? (n - tile_first) : (size_t)kTile;
    float4* dst = reinterpret_cast<float4*>(feat) + tile_first * 16;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int idx = threadIdx.x + kTile * k;
        const int row = idx >> 4, col = idx & 15;
        if ((size_t)row < rows) dst[idx] = tile[row * kPitch4 + col];
    }
}

// Optional second product of a feature pass (DecodeFuse, kernels.h): the tensor front-end's A' operand row, |x'| and the
// running max |x| of the block this thread just converted -- what col_sums / max_abs / prep_blocks otherwise re-read
// the 256 MB feature matrix for.  Same arithmetic as assign_tc.cu prep_blocks_kernel and update.cu max_abs_kernel.
// The thread's row goes through the shared-memory tile again so the 256-byte rows leave as coalesced 16-byte stores.
__device__ __forceinline__ void store_operands_tile(const float4 (&f)[16], bool valid, float4* tile, const float* s_mu, const DecodeFuse& fu,
                                                    size_t tile_first, size_t blk, size_t n) {
    __syncthreads();   // the feature tile has been streamed out
    float ss = 0.0f, mx = 0.0f;
    if (valid) {
        uint32_t* row = reinterpret_cast<uint32_t*>(tile + threadIdx.x * kPitch4);   // 64 words = 128 bf16
        auto absmax = [&](float v) {
            const float a = fabsf(v);
            if (a != a) mx = INFINITY;
            else if (a > mx) mx = a;
        };
#pragma unroll
        for (int q = 0; q < 16; ++q) {
            const float4 x = f[q];
            absmax(x.x); absmax(x.y); absmax(x.z); absmax(x.w);
            const float4 m = reinterpret_cast<const float4*>(s_mu)[q];
            const float o0 = 2.0f * (x.x - m.x), o1 = x.y - m.y, o2 = x.z - m.z, o3 = x.w - m.w;
            ss += o0 * o0 + o1 * o1 + o2 * o2 + o3 * o3;
            uint16_t h0, h1, h2, h3, l0, l1, l2, l3;
            split_bf16(o0, h0, l0);
            split_bf16(o1, h1, l1);
            split_bf16(o2, h2, l2);
            split_bf16(o3, h3, l3);
            if (fu.rgb) {
                uint16_t* r16 = reinterpret_cast<uint16_t*>(row);
                r16[3 * q] = h0; r16[3 * q + 1] = h1; r16[3 * q + 2] = h2;
                r16[64 + 3 * q] = l0; r16[64 + 3 * q + 1] = l1; r16[64 + 3 * q + 2] = l2;
                r16[48 + q] = q < 3 ? (uint16_t)0x3f80u : (uint16_t)0u;
                r16[112 + q] = 0;
            } else {
                row[2 * q] = (uint32_t)h0 | ((uint32_t)h1 << 16);
                row[2 * q + 1] = (uint32_t)h2 | ((uint32_t)h3 << 16);
                row[32 + 2 * q] = (uint32_t)l0 | ((uint32_t)l1 << 16);
                row[32 + 2 * q + 1] = (uint32_t)l2 | ((uint32_t)l3 << 16);
            }
        }
        fu.xnorm[blk] = sqrtf(ss) * 1.0001f;
    }
    __syncthreads();
    const size_t rows = (n - tile_first) < (size_t)kTile ? (n - tile_first) : (size_t)kTile;
    float4* dst = reinterpret_cast<float4*>(fu.a_prime) + tile_first * 16;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int idx = threadIdx.x + kTile * k;
        const int r = idx >> 4, col = idx & 15;
        if ((size_t)r < rows) dst[idx] = tile[r * kPitch4 + col];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) atomicMax(fu.max_abs_bits, __float_as_uint(mx));
}

// ---- K1: BCn -> pixel planes + features ----------------------------------------------------------

template <int FMT, int CH>
__global__ void __launch_bounds__(kTile) decode_bcn_kernel(const uint8_t* __restrict__ bc, size_t n,
                                                           uint4* __restrict__ planes, size_t stride,
                                                           float* __restrict__ feat, uint8_t* __restrict__ px_blocks,
                                                           int vec16, DecodeFuse fu) {
    __shared__ float4 tile[kTile * kPitch4];
    __shared__ __align__(16) float s_mu[64];
    if (fu.a_prime && threadIdx.x < 64) s_mu[threadIdx.x] = fu.mu[threadIdx.x];   // visible after the tile barrier below
    const size_t tile_first = (size_t)blockIdx.x * kTile;
    const size_t blk = tile_first + threadIdx.x;
    const bool valid = blk < n;
    uint32_t w[4 * CH];
    float4 f[16];
    constexpr bool k8 = (FMT == YUPVQ_FMT_BC1 || FMT == YUPVQ_FMT_BC4);

    uint4 raw = make_uint4(0, 0, 0, 0);
    if (k8) {
        // 8-byte blocks: even lanes fetch 16 bytes (two blocks) and hand the upper half to lane+1.
        if (vec16) {
            uint4 v = make_uint4(0, 0, 0, 0);
            const bool even = (threadIdx.x & 1) == 0;
            if (even && blk + 1 < n) v = ldg_nc_v4(reinterpret_cast<const uint4*>(bc) + (blk >> 1));
            else if (even && valid) {
                const uint2 t2 = *(reinterpret_cast<const uint2*>(bc) + blk);
                v.x = t2.x;
                v.y = t2.y;
            }
            const uint32_t hz = __shfl_up_sync(0xffffffffu, v.z, 1), hw = __shfl_up_sync(0xffffffffu, v.w, 1);
            raw.x = even ? v.x : hz;
            raw.y = even ? v.y : hw;
        } else if (valid) {
            const uint8_t* p = bc + blk * 8;
            raw.x = p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24);
            raw.y = p[4] | (p[5] << 8) | (p[6] << 16) | ((uint32_t)p[7] << 24);
        }
    } else if (valid) {
        if (vec16) raw = ldg_nc_v4(reinterpret_cast<const uint4*>(bc) + blk);
        else {
            const uint8_t* p = bc + blk * 16;
            uint32_t t[4];
            for (int i = 0; i < 4; ++i) t[i] = p[4 * i] | (p[4 * i + 1] << 8) | (p[4 * i + 2] << 16) | ((uint32_t)p[4 * i + 3] << 24);
            raw = make_uint4(t[0], t[1], t[2], t[3]);
        }
    }

    if (valid) {
        uint32_t t[16];
        uint32_t pal[4] = {0, 0, 0, 0};
        if (FMT == YUPVQ_FMT_BC1) {
            bc1_palette(raw.x, true, pal);
#pragma unroll
            for (int p = 0; p < 16; ++p) {
                const uint32_t k = (raw.y >> (2 * p)) & 3u;
                t[p] = k == 0 ? pal[0] : k == 1 ? pal[1] : k == 2 ? pal[2] : pal[3];
            }
        } else if (FMT == YUPVQ_FMT_BC3) {
            uint32_t a[16];
            bc1_palette(raw.z, false, pal);
            bc4_decode(raw.x, raw.y, a);
#pragma unroll
            for (int p = 0; p < 16; ++p) {
                const uint32_t k = (raw.w >> (2 * p)) & 3u;
                const uint32_t c = k == 0 ? pal[0] : k == 1 ? pal[1] : k == 2 ? pal[2] : pal[3];
                t[p] = (c & 0x00ffffffu) | (a[p] << 24);
            }
        } else if (FMT == YUPVQ_FMT_BC4) {
            uint32_t a[16];
            bc4_decode(raw.x, raw.y, a);
#pragma unroll
            for (int p = 0; p < 16; ++p) t[p] = pack_rgba(a[p], a[p], a[p], 255);
        } else if (FMT == YUPVQ_FMT_BC5) {
            uint32_t a[16], b[16];
            bc4_decode(raw.x, raw.y, a);
            bc4_decode(raw.z, raw.w, b);
#pragma unroll
            for (int p = 0; p < 16; ++p) t[p] = pack_rgba(a[p], b[p], 0, 255);
        } else {
            bc7_decode(raw, t);
        }
        pack_block<FMT, CH>(t, w);
        if (planes) store_planes<CH>(w, planes, stride, blk);
        if (px_blocks) {
            uint32_t* dst = reinterpret_cast<uint32_t*>(px_blocks + blk * 16 * CH);
#pragma unroll
            for (int i = 0; i < 4 * CH; ++i) dst[i] = w[i];
        }
        if (feat) {
            if (FMT == YUPVQ_FMT_BC1) palette_features<CH>(pal, raw.y, w, f);
            else if (FMT == YUPVQ_FMT_BC3) palette_features<CH>(pal, raw.w, w, f);
            else block_features<CH>(w, f);
        }
    }
    if (feat) store_features_tile(f, valid, tile, feat, tile_first, n);
    if (feat && fu.a_prime) store_operands_tile(f, valid, tile, s_mu, fu, tile_first, blk, n);
}

// ---- uint8 pixel blocks (reference layout) -> planes + features ---------------------------------

template <int CH>
__global__ void __launch_bounds__(kTile) pixels_kernel(const uint8_t* __restrict__ px_blocks, size_t n,
                                                       uint4* __restrict__ planes, size_t stride,
                                                       float* __restrict__ feat, int aligned) {
    __shared__ float4 tile[kTile * kPitch4];
    const size_t tile_first = (size_t)blockIdx.x * kTile;
    const size_t blk = tile_first + threadIdx.x;
    const bool valid = blk < n;
    uint32_t w[4 * CH];
    float4 f[16];
    if (valid) {
        if (aligned) {
            const uint4* src = reinterpret_cast<const uint4*>(px_blocks) + blk * CH;
#pragma unroll
            for (int q = 0; q < CH; ++q) {
                const uint4 v = ldg_nc_v4(src + q);
                w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
            }
        } else {
            const uint8_t* p = px_blocks + blk * 16 * CH;
#pragma unroll
            for (int i = 0; i < 4 * CH; ++i)
                w[i] = p[4 * i] | (p[4 * i + 1] << 8) | (p[4 * i + 2] << 16) | ((uint32_t)p[4 * i + 3] << 24);
        }
        if (planes) store_planes<CH>(w, planes, stride, blk);
        if (feat) block_features<CH>(w, f);
    }
    if (feat) store_features_tile(f, valid, tile, feat, tile_first, n);
}

// ---- K1h: float tiles / raw float image -> features ----------------------------------------------

// VQEncoder::RgbaFloatBlockToOklabBlock, src/vq_encoder.cpp:179-185.  The conversion (fp64 cbrt / pow) is the cost, and HDR
// blocks are often piecewise constant (skies, the synthetic skybox), so it is de-duplicated the way block_features does it
// for LDR blocks, in three steps that keep the warp converged through the fp64 code: (1) compares only: every texel gets
// the slot of its (r, g, b) bit pattern in a 4-entry key list, or an overflow bit once the block has shown a fifth colour;
// (2) slot j is converted by all lanes together while any lane of the warp has more than j colours; (3) texels pick their
// slot, overflow texels are converted one by one (blocks with many colours: no worse than a plain loop).  The r01 version
// kept a one-entry cache inside the 16-texel loop: lanes missed at different texels and the warp walked the fp64 code at
// every position where any lane missed.  Same function on the same bits: identical features.
template <int CH>
__device__ __forceinline__ void float_block_features(const float (&v)[16 * CH], float4 (&f)[16]) {
    auto rgb_of = [&](int p, float& r, float& g, float& b) {
        if (CH == 1) r = g = b = v[p];
        else if (CH == 2) r = g = b = v[2 * p];
        else { r = v[p * CH]; g = v[p * CH + 1]; b = v[p * CH + 2]; }
    };
    uint32_t kr[4], kg[4], kb[4];
    uint32_t slots = 0, overflow = 0;
    int count = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) kr[j] = kg[j] = kb[j] = 0u;
#pragma unroll
    for (int p = 0; p < 16; ++p) {
        float r, g, b;
        rgb_of(p, r, g, b);
        const uint32_t ur = __float_as_uint(r), ug = __float_as_uint(g), ub = __float_as_uint(b);
        uint32_t s = 4;
#pragma unroll
        for (int j = 3; j >= 0; --j)
            if (j < count && ur == kr[j] && ug == kg[j] && ub == kb[j]) s = (uint32_t)j;
        if (s == 4) {
            if (count < 4) {
                s = (uint32_t)count;
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (j == count) { kr[j] = ur; kg[j] = ug; kb[j] = ub; }
                ++count;
            } else {
                s = 0;
                overflow |= 1u << p;
            }
        }
        slots |= s << (2 * p);
    }
    float3 val[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        val[j] = make_float3(0.f, 0.f, 0.f);
        if (__any_sync(__activemask(), count > j)) {
            const bool has = count > j;
            val[j] = float_to_oklab(has ? __uint_as_float(kr[j]) : 0.0f, has ? __uint_as_float(kg[j]) : 0.0f, has ? __uint_as_float(kb[j]) : 0.0f);
        }
    }
#pragma unroll
    for (int p = 0; p < 16; ++p) {
        const uint32_t s = (slots >> (2 * p)) & 3u;
        float3 o = s == 0 ? val[0] : s == 1 ? val[1] : s == 2 ? val[2] : val[3];
        if (overflow & (1u << p)) {
            float r, g, b;
            rgb_of(p, r, g, b);
            o = float_to_oklab(r, g, b);
        }
        float alpha = 1.0f;
        if (CH == 2) alpha = v[2 * p + 1];
        if (CH == 4) alpha = v[p * CH + 3];
        f[p] = make_float4(o.x, o.y, o.z, alpha);
    }
}

template <int CH>
__global__ void __launch_bounds__(kTile) float_blocks_kernel(const float* __restrict__ blocks, size_t n,
                                                             float* __restrict__ feat) {
    __shared__ float4 tile[kTile * kPitch4];
    const size_t tile_first = (size_t)blockIdx.x * kTile;
    const size_t blk = tile_first + threadIdx.x;
    const bool valid = blk < n;
    float v[16 * CH];
    float4 f[16];
    if (valid) {
        const float* src = blocks + blk * 16 * CH;
#pragma unroll
        for (int i = 0; i < 16 * CH; ++i) v[i] = src[i];
        float_block_features<CH>(v, f);
    }
    store_features_tile(f, valid, tile, feat, tile_first, n);
}

// The edge-clamped 4x4 gather of src/vq_bcn_compressor.cpp:235-252 fused with the conversion.
template <int CH>
__global__ void __launch_bounds__(kTile) image_kernel(const float* __restrict__ img, uint32_t width, uint32_t height,
                                                      size_t n, float* __restrict__ feat, float* __restrict__ tiles) {
    __shared__ float4 tile[kTile * kPitch4];
    const size_t tile_first = (size_t)blockIdx.x * kTile;
    const size_t blk = tile_first + threadIdx.x;
    const bool valid = blk < n;
    const size_t bx = (width + 3) / 4;
    float v[16 * CH];
    float4 f[16];
    if (valid) {
        const size_t x0 = (blk % bx) * 4, y0 = (blk / bx) * 4;
#pragma unroll
        for (int y = 0; y < 4; ++y) {
            const size_t py = min(y0 + y, (size_t)height - 1);
#pragma unroll
            for (int x = 0; x < 4; ++x) {
                const size_t px = min(x0 + x, (size_t)width - 1);
                const float* s = img + (py * width + px) * CH;
#pragma unroll
                for (int c = 0; c < CH; ++c) v[(y * 4 + x) * CH + c] = s[c];
            }
        }
        if (tiles) {
            float* d = tiles + blk * 16 * CH;
#pragma unroll
            for (int i = 0; i < 16 * CH; ++i) d[i] = v[i];
        }
        if (feat) float_block_features<CH>(v, f);
    }
    if (feat) store_features_tile(f, valid, tile, feat, tile_first, n);
}

// ---- K6: centroids back to pixels -----------------------------------------------------------------

// VQEncoder::OklabToPixel, src/vq_encoder.cpp:67-93: one thread per texel.
__global__ void oklab_to_pixels_kernel(const float* __restrict__ oklab, size_t n_px, int ch, uint8_t* __restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_px) return;
    const float4 o = reinterpret_cast<const float4*>(oklab)[i];
    double rgb[3];
    oklab_to_rgb255(o.x, o.y, o.z, rgb);
    const unsigned r = to_u8(rgb[0]), g = to_u8(rgb[1]), b = to_u8(rgb[2]);
    const float af = fminf(fmaxf(__fmul_rn(o.w, 255.0f), 0.0f), 255.0f);
    const unsigned a = (unsigned)af;
    uint8_t* d = out + i * ch;
    switch (ch) {
        case 1: d[0] = (uint8_t)luma709((float)r, (float)g, (float)b); break;
        case 2: d[0] = (uint8_t)luma709((float)r, (float)g, (float)b); d[1] = (uint8_t)a; break;
        case 3: d[0] = (uint8_t)r; d[1] = (uint8_t)g; d[2] = (uint8_t)b; break;
        default: d[0] = (uint8_t)r; d[1] = (uint8_t)g; d[2] = (uint8_t)b; d[3] = (uint8_t)a; break;
    }
}

// VQEncoder::OklabToRgbaFloat, src/vq_encoder.cpp:150-176 (clip() is a discarded copy: no clamp).
__global__ void oklab_to_floats_kernel(const float* __restrict__ oklab, size_t n_px, int ch, float* __restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_px) return;
    const float4 o = reinterpret_cast<const float4*>(oklab)[i];
    double rgb[3];
    oklab_to_rgb255(o.x, o.y, o.z, rgb);
    const float r = (float)rgb[0], g = (float)rgb[1], b = (float)rgb[2];
    float* d = out + i * ch;
    switch (ch) {
        case 1: d[0] = luma709(r, g, b); break;
        case 2: d[0] = luma709(r, g, b); d[1] = o.w; break;
        case 3: d[0] = r; d[1] = g; d[2] = b; break;
        default: d[0] = r; d[1] = g; d[2] = b; d[3] = o.w; break;
    }
}

// ---- gathers ---------------------------------------------------------------------------------------

__global__ void gather_rows_kernel(const float* __restrict__ src, const uint32_t* __restrict__ rows, size_t n_rows,
                                   int row_f4, float* __restrict__ dst) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rows * row_f4) return;
    const size_t r = i / row_f4, c = i % row_f4;
    reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(src)[(size_t)rows[r] * row_f4 + c];
}

__global__ void gather_blocks_kernel(const uint4* __restrict__ planes, size_t stride, const float* __restrict__ feat,
                                     const uint32_t* __restrict__ rows, size_t n_rows, int ch,
                                     uint4* __restrict__ out_planes, size_t out_stride, float* __restrict__ out_feat) {
    // 16 threads per row: each moves one float4 of features; the first `ch` of them also move a plane word
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rows * 16) return;
    const size_t r = i >> 4;
    const int c = (int)(i & 15);
    const size_t src = rows[r];
    if (out_feat) reinterpret_cast<float4*>(out_feat)[i] = reinterpret_cast<const float4*>(feat)[src * 16 + c];
    if (out_planes && c < ch) out_planes[c * out_stride + r] = planes[c * stride + src];
}

__global__ void planes_to_blocks_kernel(const uint4* __restrict__ planes, size_t stride, const uint32_t* __restrict__ rows,
                                        size_t n_rows, int ch, uint8_t* __restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rows * ch) return;
    const size_t r = i / ch;
    const int q = (int)(i % ch);
    const size_t src = rows ? rows[r] : r;
    const uint4 v = planes[q * stride + src];
    uint32_t* d = reinterpret_cast<uint32_t*>(out + r * 16 * ch + q * 16);
    d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
}

// VQBCnCompressor::DecompressToBCn gather, src/vq_bcn_compressor.cpp:350-358: out[i] = codebook[indices[i]].
// HBM-bound (4 B read + blockSize B written per block): the codebook (<= 64 KB) is staged in shared memory, a thread takes
// four consecutive indices with one 16-byte load and writes 4 x blockSize contiguous bytes with streaming stores, so a warp
// covers 1-2 KB of output per step; grid = a few CTAs per SM, grid-stride.
template <typename V>
__device__ __forceinline__ void st_cs(V* p, const V& v);
template <>
__device__ __forceinline__ void st_cs<uint2>(uint2* p, const uint2& v) {
    asm volatile("st.global.cs.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}
template <>
__device__ __forceinline__ void st_cs<uint4>(uint4* p, const uint4& v) {
    asm volatile("st.global.cs.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

template <typename V>
__global__ void __launch_bounds__(256) gather_codebook_kernel(const V* __restrict__ codebook, uint32_t n_entries,
                                                              const uint32_t* __restrict__ indices, size_t n, V* __restrict__ out,
                                                              int staged) {
    extern __shared__ uint4 gather_smem[];
    const V* cb = codebook;
    if (staged) {
        V* cb_s = reinterpret_cast<V*>(gather_smem);
        for (uint32_t k = threadIdx.x; k < n_entries; k += blockDim.x) cb_s[k] = codebook[k];
        __syncthreads();
        cb = cb_s;
    }
    auto fetch = [&](uint32_t k) {   // an out-of-range index leaves the block zero (the reference leaves it default-initialised)
        V v = V();
        if (k < n_entries) v = cb[k];
        return v;
    };
    const size_t n4 = n / 4;
    const bool aligned = (reinterpret_cast<uintptr_t>(indices) & 15u) == 0;
    for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < n4; g += (size_t)gridDim.x * blockDim.x) {
        uint4 k;
        if (aligned) {
            k = ldg_nc_v4(reinterpret_cast<const uint4*>(indices) + g);
        } else {
            k = make_uint4(indices[4 * g], indices[4 * g + 1], indices[4 * g + 2], indices[4 * g + 3]);
        }
        const V v0 = fetch(k.x), v1 = fetch(k.y), v2 = fetch(k.z), v3 = fetch(k.w);
        V* o = out + 4 * g;
        st_cs(o, v0);
        st_cs(o + 1, v1);
        st_cs(o + 2, v2);
        st_cs(o + 3, v3);
    }
    if (blockIdx.x == 0)
        for (size_t i = n4 * 4 + threadIdx.x; i < n; i += blockDim.x) out[i] = fetch(indices[i]);
}

inline unsigned grid_for(size_t n, int per) { return (unsigned)((n + per - 1) / per); }

}  // namespace

cudaError_t colour_tables_init() {
    double t[256];
    for (int u = 0; u < 256; ++u) t[u] = host_srgb8_to_linear((double)u);
    return cudaMemcpyToSymbol(c_srgb8_to_linear, t, sizeof(t));
}

#define YUPVQ_DISPATCH_CH(KERNEL, ...)                 \
    switch (ch) {                                      \
        case 1: KERNEL<1> __VA_ARGS__; break;          \
        case 2: KERNEL<2> __VA_ARGS__; break;          \
        case 3: KERNEL<3> __VA_ARGS__; break;          \
        case 4: KERNEL<4> __VA_ARGS__; break;          \
        default: return cudaErrorInvalidValue;         \
    }

template <int FMT>
static cudaError_t decode_fmt(const uint8_t* bc, size_t n, int ch, uint4* planes, size_t stride, float* feat,
                              uint8_t* px, int vec16, cudaStream_t s, const DecodeFuse& fu) {
    const unsigned g = grid_for(n, kTile);
    switch (ch) {
        case 1: decode_bcn_kernel<FMT, 1><<<g, kTile, 0, s>>>(bc, n, planes, stride, feat, px, vec16, fu); break;
        case 2: decode_bcn_kernel<FMT, 2><<<g, kTile, 0, s>>>(bc, n, planes, stride, feat, px, vec16, fu); break;
        case 3: decode_bcn_kernel<FMT, 3><<<g, kTile, 0, s>>>(bc, n, planes, stride, feat, px, vec16, fu); break;
        case 4: decode_bcn_kernel<FMT, 4><<<g, kTile, 0, s>>>(bc, n, planes, stride, feat, px, vec16, fu); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

cudaError_t launch_decode_bcn(const uint8_t* bc, size_t n, int fmt, int ch, uint4* planes, size_t plane_stride,
                              float* feat, uint8_t* px_blocks, cudaStream_t s, const DecodeFuse* fuse) {
    if (n == 0) return cudaSuccess;
    const int vec16 = ((reinterpret_cast<uintptr_t>(bc) & 15u) == 0) ? 1 : 0;
    const DecodeFuse fu = fuse ? *fuse : DecodeFuse{};
    switch (fmt) {
        case YUPVQ_FMT_BC1: return decode_fmt<YUPVQ_FMT_BC1>(bc, n, ch, planes, plane_stride, feat, px_blocks, vec16, s, fu);
        case YUPVQ_FMT_BC3: return decode_fmt<YUPVQ_FMT_BC3>(bc, n, ch, planes, plane_stride, feat, px_blocks, vec16, s, fu);
        case YUPVQ_FMT_BC4: return decode_fmt<YUPVQ_FMT_BC4>(bc, n, ch, planes, plane_stride, feat, px_blocks, vec16, s, fu);
        case YUPVQ_FMT_BC5: return decode_fmt<YUPVQ_FMT_BC5>(bc, n, ch, planes, plane_stride, feat, px_blocks, vec16, s, fu);
        case YUPVQ_FMT_BC7: return decode_fmt<YUPVQ_FMT_BC7>(bc, n, ch, planes, plane_stride, feat, px_blocks, vec16, s, fu);
        default: return cudaErrorInvalidValue;
    }
}

// BC1 blocks -> the compact form the K-Means++ seeding keeps resident in shared memory (seed.cu, PAL variant): five
// 32-bit planes per block -- the four palette entries of R, of G and of B (byte j = entry j), and the sixteen 2-bit
// indices widened to PRMT selector nibbles (texels 0-7, texels 8-15).  20 bytes describe the same 48 pixel bytes the
// planes hold (3-channel output of a BC1 block: R, G, B of the decoded texel; entry 3 of 3-colour mode is black).
__global__ void __launch_bounds__(256) bc1_palette_planes_kernel(const uint8_t* __restrict__ bc, size_t n,
                                                                 uint32_t* __restrict__ pal5, size_t stride, int aligned) {
    const size_t blk = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (blk >= n) return;
    uint32_t lo, idx;
    if (aligned) {
        const uint2 v = __ldg(reinterpret_cast<const uint2*>(bc) + blk);
        lo = v.x;
        idx = v.y;
    } else {
        const uint8_t* p = bc + blk * 8;
        lo = p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24);
        idx = p[4] | (p[5] << 8) | (p[6] << 16) | ((uint32_t)p[7] << 24);
    }
    uint32_t pal[4];
    bc1_palette(lo, true, pal);
#pragma unroll
    for (int c = 0; c < 3; ++c)
        pal5[c * stride + blk] = ((pal[0] >> (8 * c)) & 255u) | (((pal[1] >> (8 * c)) & 255u) << 8) |
                                 (((pal[2] >> (8 * c)) & 255u) << 16) | (((pal[3] >> (8 * c)) & 255u) << 24);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        uint32_t x = (idx >> (16 * h)) & 0xffffu;  // 8 two-bit indices -> 8 nibbles
        x = (x | (x << 8)) & 0x00ff00ffu;
        x = (x | (x << 4)) & 0x0f0f0f0fu;
        x = (x | (x << 2)) & 0x33333333u;
        pal5[(3 + h) * stride + blk] = x;
    }
}

cudaError_t launch_bc1_palette_planes(const uint8_t* bc, size_t n, uint32_t* pal5, size_t stride, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    const int aligned = ((reinterpret_cast<uintptr_t>(bc) & 7u) == 0) ? 1 : 0;
    bc1_palette_planes_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(bc, n, pal5, stride, aligned);
    return cudaGetLastError();
}

cudaError_t launch_pixels_to_features(const uint8_t* px_blocks, size_t n, int ch, uint4* planes, size_t plane_stride,
                                      float* feat, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    const int aligned = ((reinterpret_cast<uintptr_t>(px_blocks) & 15u) == 0) ? 1 : 0;
    const unsigned g = grid_for(n, kTile);
    YUPVQ_DISPATCH_CH(pixels_kernel, <<<g, kTile, 0, s>>>(px_blocks, n, planes, plane_stride, feat, aligned))
    return cudaGetLastError();
}

cudaError_t launch_floats_to_features(const float* float_blocks, size_t n, int ch, float* feat, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    const unsigned g = grid_for(n, kTile);
    YUPVQ_DISPATCH_CH(float_blocks_kernel, <<<g, kTile, 0, s>>>(float_blocks, n, feat))
    return cudaGetLastError();
}

cudaError_t launch_image_to_features(const float* image, uint32_t w, uint32_t h, int ch, float* feat, float* tiles,
                                     cudaStream_t s) {
    const size_t n = (size_t)((w + 3) / 4) * ((h + 3) / 4);
    if (n == 0) return cudaSuccess;
    const unsigned g = grid_for(n, kTile);
    YUPVQ_DISPATCH_CH(image_kernel, <<<g, kTile, 0, s>>>(image, w, h, n, feat, tiles))
    return cudaGetLastError();
}

cudaError_t launch_oklab_to_pixels(const float* oklab, size_t n, int ch, uint8_t* out, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    oklab_to_pixels_kernel<<<grid_for(n * 16, 128), 128, 0, s>>>(oklab, n * 16, ch, out);
    return cudaGetLastError();
}

cudaError_t launch_oklab_to_floats(const float* oklab, size_t n, int ch, float* out, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    oklab_to_floats_kernel<<<grid_for(n * 16, 128), 128, 0, s>>>(oklab, n * 16, ch, out);
    return cudaGetLastError();
}

cudaError_t launch_gather_rows_f32(const float* src, const uint32_t* rows, size_t n_rows, int row_floats, float* dst,
                                   cudaStream_t s) {
    if (n_rows == 0) return cudaSuccess;
    const int f4 = row_floats / 4;
    gather_rows_kernel<<<grid_for(n_rows * f4, 256), 256, 0, s>>>(src, rows, n_rows, f4, dst);
    return cudaGetLastError();
}

cudaError_t launch_gather_blocks(const uint4* planes, size_t plane_stride, const float* feat, const uint32_t* rows,
                                 size_t n_rows, int ch, uint4* out_planes, size_t out_stride, float* out_feat,
                                 cudaStream_t s) {
    if (n_rows == 0) return cudaSuccess;
    gather_blocks_kernel<<<grid_for(n_rows * 16, 256), 256, 0, s>>>(planes, plane_stride, feat, rows, n_rows, ch,
                                                                    out_planes, out_stride, out_feat);
    return cudaGetLastError();
}

cudaError_t launch_planes_to_blocks(const uint4* planes, size_t plane_stride, const uint32_t* rows, size_t n_rows,
                                    int ch, uint8_t* out_px_blocks, cudaStream_t s) {
    if (n_rows == 0) return cudaSuccess;
    planes_to_blocks_kernel<<<grid_for(n_rows * ch, 256), 256, 0, s>>>(planes, plane_stride, rows, n_rows, ch,
                                                                       out_px_blocks);
    return cudaGetLastError();
}

cudaError_t launch_gather_codebook(const uint8_t* codebook, uint32_t n_entries, uint32_t block_size,
                                   const uint32_t* indices, size_t n, uint8_t* out, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    if (block_size != 8 && block_size != 16) return cudaErrorInvalidValue;
    const size_t cb_bytes = (size_t)n_entries * block_size;
    const int staged = cb_bytes <= 48 * 1024 ? 1 : 0;   // the reference's K <= 1024 -> <= 16 KB
    const size_t smem = staged ? cb_bytes : 0;
    size_t groups = (n / 4 + 255) / 256;
    unsigned grid = (unsigned)std::min<size_t>(std::max<size_t>(groups, 1), (size_t)YUPVQ_NUM_SMS * 8);
    if (block_size == 8)
        gather_codebook_kernel<uint2><<<grid, 256, smem, s>>>(reinterpret_cast<const uint2*>(codebook), n_entries, indices, n,
                                                              reinterpret_cast<uint2*>(out), staged);
    else
        gather_codebook_kernel<uint4><<<grid, 256, smem, s>>>(reinterpret_cast<const uint4*>(codebook), n_entries, indices, n,
                                                              reinterpret_cast<uint4*>(out), staged);
    return cudaGetLastError();
}

}  // namespace yupvq
