// Device BCn block decoders (BC1, BC3, BC4, BC5, BC7) producing 16 RGBA8 texels per block, bit-exact
// with the reference's decoder: BCnCompressor::Decompress -> Compressonator 4.3 CMP_ConvertTexture
// (reference src/bcn_compressor.cpp:191-227, called per block at src/vq_bcn_compressor.cpp:142-146).
// Observed rules are listed in SURVEY.md Appendix A.3 and pinned by tests/golden/bc_decode.npz.
#pragma once
#include "common.cuh"
// tables live in global memory (L1-cached): one 16-byte row load per block; divergent indices
// would serialise in the constant cache.
#define YUPVQ_TABLE_QUAL __device__ __align__(16)
#include "bc7_tables.h"

namespace yupvq {

__device__ __forceinline__ uint32_t pack_rgba(int r, int g, int b, int a) {
    return (uint32_t)r | ((uint32_t)g << 8) | ((uint32_t)b << 16) | ((uint32_t)a << 24);
}

// RGB565 endpoint expansion by bit replication.
__device__ __forceinline__ void rgb565(uint32_t c, int& r, int& g, int& b) {
    const int r5 = (c >> 11) & 31, g6 = (c >> 5) & 63, b5 = c & 31;
    r = (r5 << 3) | (r5 >> 2);
    g = (g6 << 2) | (g6 >> 4);
    b = (b5 << 3) | (b5 >> 2);
}

// BC1 colour block (lo = endpoints, hi = 2-bit selectors) -> 4-entry RGBA palette.
// allow3: BC1 switches to 3-colour + transparent black when c0 <= c1; BC3's colour half never does.
__device__ __forceinline__ void bc1_palette(uint32_t lo, bool allow3, uint32_t pal[4]) {
    const uint32_t c0 = lo & 0xffffu, c1 = lo >> 16;
    int r0, g0, b0, r1, g1, b1;
    rgb565(c0, r0, g0, b0);
    rgb565(c1, r1, g1, b1);
    pal[0] = pack_rgba(r0, g0, b0, 255);
    pal[1] = pack_rgba(r1, g1, b1, 255);
    if (c0 > c1 || !allow3) {
        pal[2] = pack_rgba((2 * r0 + r1) / 3, (2 * g0 + g1) / 3, (2 * b0 + b1) / 3, 255);
        pal[3] = pack_rgba((r0 + 2 * r1) / 3, (g0 + 2 * g1) / 3, (b0 + 2 * b1) / 3, 255);
    } else {
        pal[2] = pack_rgba((r0 + r1) / 2, (g0 + g1) / 2, (b0 + b1) / 2, 255);
        pal[3] = 0u;
    }
}

// BC3 alpha / BC4 / BC5 channel block -> 8-entry palette.
__device__ __forceinline__ void bc4_palette(uint32_t a0, uint32_t a1, uint32_t pal[8]) {
    pal[0] = a0;
    pal[1] = a1;
    if (a0 > a1) {
#pragma unroll
        for (int i = 2; i < 8; ++i) pal[i] = ((8 - i) * a0 + (i - 1) * a1 + 3) / 7;
    } else {
#pragma unroll
        for (int i = 2; i < 6; ++i) pal[i] = ((6 - i) * a0 + (i - 1) * a1 + 2) / 5;
        pal[6] = 0;
        pal[7] = 255;
    }
}

__device__ __forceinline__ void bc4_decode(uint32_t w0, uint32_t w1, uint32_t out[16]) {
    uint32_t pal[8];
    bc4_palette(w0 & 255u, (w0 >> 8) & 255u, pal);
    const unsigned long long bits = ((unsigned long long)w1 << 16) | (w0 >> 16);
#pragma unroll
    for (int p = 0; p < 16; ++p) {
        const uint32_t k = (uint32_t)(bits >> (3 * p)) & 7u;
        uint32_t v = pal[0];
#pragma unroll
        for (int j = 1; j < 8; ++j) v = (k == (uint32_t)j) ? pal[j] : v;
        out[p] = v;
    }
}

struct Bc7Bits {
    unsigned long long lo, hi;
    int pos;
    __device__ __forceinline__ uint32_t get(int n) {
        if (n == 0) return 0;
        unsigned long long v;
        if (pos >= 64) v = hi >> (pos - 64);
        else if (pos == 0) v = lo;
        else v = (lo >> pos) | (hi << (64 - pos));
        pos += n;
        return (uint32_t)(v & ((1ull << n) - 1ull));
    }
};

__device__ __forceinline__ int bc7_weight(int bits, int i) {
    // 2-bit {0,21,43,64}; 3-bit {0,9,18,27,37,46,55,64}; 4-bit {0,4,9,13,17,21,26,30,34,38,43,47,51,55,60,64}
    if (bits == 2) return (i * 64 + 1) / 3;                 // 0,21,43,64
    if (bits == 3) return (i * 64 + 3) / 7;                 // 0,9,18,27,37,46,55,64
    return (i * 64 + 7) / 15;                               // 4-bit table
}

__device__ __forceinline__ int bc7_interp(int a, int b, int w) { return (a * (64 - w) + b * w + 32) >> 6; }

// BC7 block -> 16 RGBA8 texels.  The reserved encoding (first byte 0) decodes to zeros here; the
// reference leaves stale memory for it and no encoder emits it.
__device__ void bc7_decode(uint4 blk, uint32_t out[16]) {
    const uint32_t b0 = blk.x & 255u;
    if (b0 == 0) {
#pragma unroll
        for (int p = 0; p < 16; ++p) out[p] = 0;
        return;
    }
    const int mode = __ffs((int)b0) - 1;
    // per-mode parameters packed as nibbles (mode 0 in the low nibble)
    const int ns = (0x21112323u >> (4 * mode)) & 15;
    const int pb = (0x60006664u >> (4 * mode)) & 15;
    const int rb = (0x00220000u >> (4 * mode)) & 15;
    const int isb = (0x00010000u >> (4 * mode)) & 15;
    const int cb = (0x57757564u >> (4 * mode)) & 15;
    const int ab = (0x57860000u >> (4 * mode)) & 15;
    const int epb = (0x11001001u >> (4 * mode)) & 15;
    const int spb = (0x00000010u >> (4 * mode)) & 15;
    const int ib = (0x24222233u >> (4 * mode)) & 15;
    const int ib2 = (0x00230000u >> (4 * mode)) & 15;

    Bc7Bits br;
    br.lo = ((unsigned long long)blk.y << 32) | blk.x;
    br.hi = ((unsigned long long)blk.w << 32) | blk.z;
    br.pos = mode + 1;
    const int part = (int)br.get(pb);
    const int rot = (int)br.get(rb);
    const int isel = (int)br.get(isb);
    int ep[6][4];
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int e = 0; e < 6; ++e)
            if (e < 2 * ns) ep[e][c] = (int)br.get(cb);
#pragma unroll
    for (int e = 0; e < 6; ++e)
        if (e < 2 * ns) ep[e][3] = ab ? (int)br.get(ab) : 255;
    int cbits = cb, abits = ab;
    if (epb) {
#pragma unroll
        for (int e = 0; e < 6; ++e)
            if (e < 2 * ns) {
                const int p = (int)br.get(1);
                ep[e][0] = (ep[e][0] << 1) | p;
                ep[e][1] = (ep[e][1] << 1) | p;
                ep[e][2] = (ep[e][2] << 1) | p;
                if (abits) ep[e][3] = (ep[e][3] << 1) | p;
            }
        ++cbits;
        if (abits) ++abits;
    } else if (spb) {
#pragma unroll
        for (int s = 0; s < 3; ++s)
            if (s < ns) {
                const int p = (int)br.get(1);
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    ep[2 * s][c] = (ep[2 * s][c] << 1) | p;
                    ep[2 * s + 1][c] = (ep[2 * s + 1][c] << 1) | p;
                }
            }
        ++cbits;
    }
#pragma unroll
    for (int e = 0; e < 6; ++e)
        if (e < 2 * ns) {
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const int v = ep[e][c] << (8 - cbits);
                ep[e][c] = v | (v >> cbits);
            }
            if (abits) {
                const int v = ep[e][3] << (8 - abits);
                ep[e][3] = v | (v >> abits);
            }
        }
    int anchor1 = -1, anchor2 = -1;
    if (ns == 2) anchor1 = kBc7Anchor2[part];
    if (ns == 3) {
        anchor1 = kBc7Anchor3a[part];
        anchor2 = kBc7Anchor3b[part];
    }
    uint32_t sub = 0;  // 2 bits per texel
    if (ns > 1) {
        const uint4 row = *reinterpret_cast<const uint4*>(ns == 2 ? kBc7Partition2[part] : kBc7Partition3[part]);
        const uint32_t w[4] = {row.x, row.y, row.z, row.w};
#pragma unroll
        for (int p = 0; p < 16; ++p) sub |= ((w[p >> 2] >> (8 * (p & 3))) & 3u) << (2 * p);
    }
    unsigned long long idx1 = 0, idx2 = 0;  // 4 bits per texel
#pragma unroll
    for (int p = 0; p < 16; ++p) {
        const bool is_anchor = (p == 0) || (p == anchor1) || (p == anchor2);
        idx1 |= (unsigned long long)br.get(ib - (is_anchor ? 1 : 0)) << (4 * p);
    }
    if (ib2) {
#pragma unroll
        for (int p = 0; p < 16; ++p) idx2 |= (unsigned long long)br.get(ib2 - (p == 0 ? 1 : 0)) << (4 * p);
    }
#pragma unroll
    for (int p = 0; p < 16; ++p) {
        const int s = (int)((sub >> (2 * p)) & 3u);
        int e0[4], e1[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            e0[c] = s == 0 ? ep[0][c] : s == 1 ? ep[2][c] : ep[4][c];
            e1[c] = s == 0 ? ep[1][c] : s == 1 ? ep[3][c] : ep[5][c];
        }
        const int i1 = (int)((idx1 >> (4 * p)) & 15u), i2 = (int)((idx2 >> (4 * p)) & 15u);
        int ci = i1, ai = i1, cib = ib, aib = ib;
        if (ib2) {
            if (isel) { ci = i2; cib = ib2; }
            else { ai = i2; aib = ib2; }
        }
        const int wc = bc7_weight(cib, ci), wa = bc7_weight(aib, ai);
        int r = bc7_interp(e0[0], e1[0], wc), g = bc7_interp(e0[1], e1[1], wc), b = bc7_interp(e0[2], e1[2], wc);
        int a = bc7_interp(e0[3], e1[3], wa);
        int t;
        if (rot == 1) { t = r; r = a; a = t; }
        if (rot == 2) { t = g; g = a; a = t; }
        if (rot == 3) { t = b; b = a; a = t; }
        out[p] = pack_rgba(r, g, b, a);
    }
}

}  // namespace yupvq
