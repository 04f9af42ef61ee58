// Block encoders of bcn_encode.cu (BC1 colour blocks, BC4-style scalar blocks) as plain functions of 16 texels.
// Device code in the product; the same source also compiles as host C++ (YUPVQ_HD empty) for the CPU-side check in
// tests/cpp/bcn_encode_host.cpp, which lets the -m "not gpu" suite gate the encoders' PSNR against Compressonator.
#pragma once
#include <stdint.h>
#if defined(__CUDACC__)
#define YUPVQ_HD __device__ __forceinline__
#define YUPVQ_HD_NOINLINE __device__
#else
#include <algorithm>
#include <cmath>
#define YUPVQ_HD inline
#define YUPVQ_HD_NOINLINE inline
#endif

namespace yupvq {
namespace bcnenc {

#if defined(__CUDACC__)
YUPVQ_HD int f2i_rn(float x) { return __float2int_rn(x); }
YUPVQ_HD int popc32(uint32_t x) { return __popc(x); }
YUPVQ_HD int imin(int a, int b) { return min(a, b); }
YUPVQ_HD int imax(int a, int b) { return max(a, b); }
struct U2 { uint32_t x, y; };
#else
inline int f2i_rn(float x) { return (int)std::nearbyintf(x); }
inline int popc32(uint32_t x) { return __builtin_popcount(x); }
inline int imin(int a, int b) { return a < b ? a : b; }
inline int imax(int a, int b) { return a > b ? a : b; }
struct U2 { uint32_t x, y; };
#endif

#if defined(__CUDACC__)
#define YUPVQ_ENC_TABLE static __device__ const
#else
#define YUPVQ_ENC_TABLE static const
#endif
#include "bc1_single_tables.h"

YUPVQ_HD uint32_t pack565f(float r, float g, float b) {
    const int r5 = imin(31, imax(0, f2i_rn(r * (31.0f / 255.0f))));
    const int g6 = imin(63, imax(0, f2i_rn(g * (63.0f / 255.0f))));
    const int b5 = imin(31, imax(0, f2i_rn(b * (31.0f / 255.0f))));
    return (uint32_t)((r5 << 11) | (g6 << 5) | b5);
}

YUPVQ_HD void expand565(uint32_t c, int& r, int& g, int& b) {
    const int r5 = (c >> 11) & 31, g6 = (c >> 5) & 63, b5 = c & 31;
    r = (r5 << 3) | (r5 >> 2);
    g = (g6 << 2) | (g6 >> 4);
    b = (b5 << 3) | (b5 >> 2);
}

// Selectors and squared error of one endpoint pair.  four: palette {c0, c1, (2c0+c1)/3, (c0+2c1)/3}; else the 3-colour
// mode {c0, c1, (c0+c1)/2} with `transp` texels on entry 3.  Also returns the normal-equation sums of the
// least-squares refit (weights in thirds / halves of c0).
struct ColourFit {
    uint32_t sel;
    uint32_t err;
    float a00, a01, a11, b0[3], b1[3];
};

YUPVQ_HD ColourFit score_colour(const uint32_t (&px)[16], uint32_t transp, uint32_t c0, uint32_t c1, bool four) {
    int pr[4], pg[4], pb[4];
    expand565(c0, pr[0], pg[0], pb[0]);
    expand565(c1, pr[1], pg[1], pb[1]);
    if (four) {
        pr[2] = (2 * pr[0] + pr[1]) / 3; pg[2] = (2 * pg[0] + pg[1]) / 3; pb[2] = (2 * pb[0] + pb[1]) / 3;
        pr[3] = (pr[0] + 2 * pr[1]) / 3; pg[3] = (pg[0] + 2 * pg[1]) / 3; pb[3] = (pb[0] + 2 * pb[1]) / 3;
    } else {
        pr[2] = (pr[0] + pr[1]) / 2; pg[2] = (pg[0] + pg[1]) / 2; pb[2] = (pb[0] + pb[1]) / 2;
        pr[3] = pr[2]; pg[3] = pg[2]; pb[3] = pb[2];   // never chosen (ties go to the lower entry)
    }
    ColourFit f;
    f.sel = 0;
    f.err = 0;
    f.a00 = f.a01 = f.a11 = 0.0f;
#pragma unroll
    for (int c = 0; c < 3; ++c) f.b0[c] = f.b1[c] = 0.0f;
    // weight of c0 per entry, in sixths: 4-colour {6, 0, 4, 2}, 3-colour {6, 0, 3}
#pragma unroll
    for (int p = 0; p < 16; ++p) {
        if ((transp >> p) & 1u) {
            f.sel |= 3u << (2 * p);
            continue;
        }
        const int r = (int)(px[p] & 255u), g = (int)((px[p] >> 8) & 255u), b = (int)((px[p] >> 16) & 255u);
        uint32_t best = 0, bd = 0xffffffffu;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (!four && k == 3) break;
            const int dr = r - pr[k], dg = g - pg[k], db = b - pb[k];
            const uint32_t d = (uint32_t)(dr * dr + dg * dg + db * db);
            if (d < bd) {
                bd = d;
                best = (uint32_t)k;
            }
        }
        f.sel |= best << (2 * p);
        f.err += bd;
        const float w0 = best == 0 ? 1.0f : (best == 1 ? 0.0f : (four ? (best == 2 ? (2.0f / 3.0f) : (1.0f / 3.0f)) : 0.5f));
        const float w1 = 1.0f - w0;
        f.a00 += w0 * w0;
        f.a01 += w0 * w1;
        f.a11 += w1 * w1;
        f.b0[0] += w0 * (float)r; f.b0[1] += w0 * (float)g; f.b0[2] += w0 * (float)b;
        f.b1[0] += w1 * (float)r; f.b1[1] += w1 * (float)g; f.b1[2] += w1 * (float)b;
    }
    return f;
}

// px: 16 texels as packed RGBA8.  allow3: BC1 (the 3-colour mode exists); punch + alpha below the threshold -> transparent.
YUPVQ_HD_NOINLINE U2 encode_colour_block(const uint32_t (&px)[16], bool allow3, bool punch, uint32_t alpha_threshold) {
    uint32_t transp = 0;
    if (allow3 && punch) {
#pragma unroll
        for (int p = 0; p < 16; ++p)
            if ((px[p] >> 24) < alpha_threshold) transp |= 1u << p;
    }
    const bool four = transp == 0;
    const int n_opaque = 16 - popc32(transp);
    if (n_opaque == 0) return U2{0u, 0xffffffffu};   // c0 == c1 == 0: every texel on the transparent entry
    // mean and covariance of the opaque texels
    float mr = 0.f, mg = 0.f, mb = 0.f;
#pragma unroll
    for (int p = 0; p < 16; ++p) {
        if ((transp >> p) & 1u) continue;
        mr += (float)(px[p] & 255u);
        mg += (float)((px[p] >> 8) & 255u);
        mb += (float)((px[p] >> 16) & 255u);
    }
    const float inv = 1.0f / (float)n_opaque;
    mr *= inv; mg *= inv; mb *= inv;
    float crr = 0.f, crg = 0.f, crb = 0.f, cgg = 0.f, cgb = 0.f, cbb = 0.f;
#pragma unroll
    for (int p = 0; p < 16; ++p) {
        if ((transp >> p) & 1u) continue;
        const float dr = (float)(px[p] & 255u) - mr, dg = (float)((px[p] >> 8) & 255u) - mg, db = (float)((px[p] >> 16) & 255u) - mb;
        crr += dr * dr; crg += dr * dg; crb += dr * db; cgg += dg * dg; cgb += dg * db; cbb += db * db;
    }
    // principal axis: start from the covariance row with the largest diagonal, four power iterations
    float vr, vg, vb;
    if (crr >= cgg && crr >= cbb) { vr = crr; vg = crg; vb = crb; }
    else if (cgg >= cbb) { vr = crg; vg = cgg; vb = cgb; }
    else { vr = crb; vg = cgb; vb = cbb; }
#pragma unroll
    for (int it = 0; it < 4; ++it) {
        const float nr = crr * vr + crg * vg + crb * vb, ng = crg * vr + cgg * vg + cgb * vb, nb = crb * vr + cgb * vg + cbb * vb;
        const float s = fmaxf(fmaxf(fabsf(nr), fabsf(ng)), fabsf(nb));
        const float k = s > 0.f ? 1.0f / s : 0.f;
        vr = nr * k; vg = ng * k; vb = nb * k;
    }
    const float vv = vr * vr + vg * vg + vb * vb;
    if (four && crr + cgg + cbb == 0.f) {
        // flat block: the endpoint pair whose 2/3 entry is nearest to the colour, per channel (bc1_single_tables.h)
        const uint32_t r = px[0] & 255u, g = (px[0] >> 8) & 255u, b = (px[0] >> 16) & 255u;
        uint32_t c0 = ((uint32_t)kSingle5[r][0] << 11) | ((uint32_t)kSingle6[g][0] << 5) | (uint32_t)kSingle5[b][0];
        uint32_t c1 = ((uint32_t)kSingle5[r][1] << 11) | ((uint32_t)kSingle6[g][1] << 5) | (uint32_t)kSingle5[b][1];
        uint32_t sel = 0xaaaaaaaau;          // entry 2 = (2 c0 + c1) / 3
        if (c0 == c1) {
            sel = 0;                         // exactly representable
        } else if (allow3 && c0 < c1) {      // BC1 needs c0 > c1 for the 4-colour mode: swapped, the same value is entry 3
            const uint32_t t = c0; c0 = c1; c1 = t;
            sel = 0xffffffffu;
        }
        return U2{c0 | (c1 << 16), sel};
    }
    float tmin = 0.f, tmax = 0.f;
    if (vv > 0.f) {
        tmin = 3.4e38f; tmax = -3.4e38f;
        const float k = 1.0f / vv;
#pragma unroll
        for (int p = 0; p < 16; ++p) {
            if ((transp >> p) & 1u) continue;
            const float t = (((float)(px[p] & 255u) - mr) * vr + ((float)((px[p] >> 8) & 255u) - mg) * vg + ((float)((px[p] >> 16) & 255u) - mb) * vb) * k;
            tmin = fminf(tmin, t);
            tmax = fmaxf(tmax, t);
        }
    }
    float hr = mr + vr * tmax, hg = mg + vg * tmax, hb = mb + vb * tmax;
    float lr = mr + vr * tmin, lg = mg + vg * tmin, lb = mb + vb * tmin;
    uint32_t best_c0 = 0, best_c1 = 0, best_sel = 0, best_err = 0xffffffffu;
#pragma unroll 1
    for (int it = 0; it < 3; ++it) {
        const uint32_t c0 = pack565f(hr, hg, hb), c1 = pack565f(lr, lg, lb);
        const ColourFit f = score_colour(px, transp, c0, c1, four);
        if (f.err < best_err) {
            best_err = f.err; best_c0 = c0; best_c1 = c1; best_sel = f.sel;
        }
        if (f.err == 0) break;
        const float det = f.a00 * f.a11 - f.a01 * f.a01;
        if (fabsf(det) < 1e-6f) break;   // all texels on one entry (or symmetric): nothing to refit
        const float k = 1.0f / det;
        hr = fminf(255.f, fmaxf(0.f, (f.a11 * f.b0[0] - f.a01 * f.b1[0]) * k));
        hg = fminf(255.f, fmaxf(0.f, (f.a11 * f.b0[1] - f.a01 * f.b1[1]) * k));
        hb = fminf(255.f, fmaxf(0.f, (f.a11 * f.b0[2] - f.a01 * f.b1[2]) * k));
        lr = fminf(255.f, fmaxf(0.f, (f.a00 * f.b1[0] - f.a01 * f.b0[0]) * k));
        lg = fminf(255.f, fmaxf(0.f, (f.a00 * f.b1[1] - f.a01 * f.b0[1]) * k));
        lb = fminf(255.f, fmaxf(0.f, (f.a00 * f.b1[2] - f.a01 * f.b0[2]) * k));
    }
    // mode by endpoint order (BC1 only): c0 > c1 selects the 4-colour mode, c0 <= c1 the 3-colour mode.  Swapping the
    // endpoints swaps entries 0 <-> 1 (and 2 <-> 3 in the 4-colour mode).
    uint32_t c0 = best_c0, c1 = best_c1, sel = best_sel;
    if (four) {
        if (c0 == c1) {
            sel = 0;                       // BC1 decodes c0 == c1 in the 3-colour mode: entry 0 is the colour either way
        } else if (allow3 && c0 < c1) {
            const uint32_t t = c0; c0 = c1; c1 = t;
            sel ^= 0x55555555u;            // 0 <-> 1, 2 <-> 3
        }
    } else if (c0 > c1) {
        const uint32_t t = c0; c0 = c1; c1 = t;
        // entries 0 <-> 1; 2 (the midpoint) and 3 (transparent) stay
        const uint32_t low = ~(sel >> 1) & 0x55555555u;   // texels on entry 0 or 1
        sel ^= low;
    }
    return U2{c0 | (c1 << 16), sel};
}

// ---- scalar blocks (BC4 and the like) ---------------------------------------------------------------------------
YUPVQ_HD void scalar_palette(int a0, int a1, int (&pal)[8]) {
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

YUPVQ_HD uint32_t scalar_error(const int (&v)[16], int a0, int a1) {
    int pal[8];
    scalar_palette(a0, a1, pal);
    uint32_t err = 0;
#pragma unroll
    for (int p = 0; p < 16; ++p) {
        int bd = 1 << 20;
#pragma unroll
        for (int k = 0; k < 8; ++k) bd = imin(bd, abs(v[p] - pal[k]));
        err += (uint32_t)(bd * bd);
    }
    return err;
}

// best endpoints of one mode: eight = the 8-value mode (a0 > a1), else the 6-value mode (a0 <= a1; texels equal to 0 /
// 255 ride on the explicit entries and stay out of the fit)
YUPVQ_HD_NOINLINE void scalar_fit(const int (&v)[16], bool eight, int& best_a0, int& best_a1, uint32_t& best_err) {
    int hi = -1, lo = 256;
#pragma unroll
    for (int p = 0; p < 16; ++p) {
        if (!eight && (v[p] == 0 || v[p] == 255)) continue;
        hi = imax(hi, v[p]);
        lo = imin(lo, v[p]);
    }
    if (hi < 0) { hi = 255; lo = 0; }   // 6-value mode with nothing in between
    auto consider = [&](int e0, int e1) {
        e0 = imin(255, imax(0, e0));
        e1 = imin(255, imax(0, e1));
        const int a0 = eight ? imax(e0, e1) : imin(e0, e1), a1 = eight ? imin(e0, e1) : imax(e0, e1);
        if (eight && a0 == a1) return;      // would decode in the other mode
        const uint32_t e = scalar_error(v, a0, a1);
        if (e < best_err) { best_err = e; best_a0 = a0; best_a1 = a1; }
    };
    int c0 = hi, c1 = lo;   // c0 = the "high" end
    {   // one least-squares refit from the selectors of (hi, lo)
        consider(hi, lo);
        int pal[8];
        const int a0 = eight ? hi : lo, a1 = eight ? lo : hi;
        scalar_palette(a0, a1, pal);
        float a00 = 0.f, a01 = 0.f, a11 = 0.f, b0 = 0.f, b1 = 0.f;
        const float steps = eight ? 7.0f : 5.0f;
#pragma unroll
        for (int p = 0; p < 16; ++p) {
            int bd = 1 << 20, bk = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int d = abs(v[p] - pal[k]);
                if (d < bd) { bd = d; bk = k; }
            }
            if (!eight && bk >= 6) continue;
            // position of entry bk on the a0 -> a1 ramp
            const float pos = bk == 0 ? 0.0f : (bk == 1 ? steps : (float)(bk - 1));
            const float w1 = pos / steps, w0 = 1.0f - w1;
            a00 += w0 * w0; a01 += w0 * w1; a11 += w1 * w1;
            b0 += w0 * (float)v[p]; b1 += w1 * (float)v[p];
        }
        const float det = a00 * a11 - a01 * a01;
        if (fabsf(det) > 1e-6f) {
            const int e0 = f2i_rn((a11 * b0 - a01 * b1) / det), e1 = f2i_rn((a00 * b1 - a01 * b0) / det);
            // e0 is the fitted a0, e1 the fitted a1
            c0 = eight ? e0 : e1;
            c1 = eight ? e1 : e0;
            c0 = imin(255, imax(0, c0));
            c1 = imin(255, imax(0, c1));
        }
    }
#pragma unroll 1
    for (int d0 = -2; d0 <= 2; ++d0)
#pragma unroll 1
        for (int d1 = -2; d1 <= 2; ++d1) consider(c0 + d0, c1 + d1);
}

YUPVQ_HD_NOINLINE U2 encode_scalar_block(const int (&v)[16]) {
    int hi = 0, lo = 255;
    bool extreme = false;
#pragma unroll
    for (int p = 0; p < 16; ++p) {
        hi = imax(hi, v[p]);
        lo = imin(lo, v[p]);
        extreme |= v[p] == 0 || v[p] == 255;
    }
    int a0 = hi, a1 = lo;
    if (hi != lo) {
        uint32_t err = 0xffffffffu;
        scalar_fit(v, true, a0, a1, err);
        if (extreme && err != 0) scalar_fit(v, false, a0, a1, err);
    }
    int pal[8];
    scalar_palette(a0, a1, pal);
    unsigned long long bits = 0;
#pragma unroll
    for (int p = 0; p < 16; ++p) {
        int bd = 1 << 20, bk = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int d = abs(v[p] - pal[k]);
            if (d < bd) { bd = d; bk = k; }
        }
        bits |= (unsigned long long)bk << (3 * p);
    }
    return U2{(uint32_t)a0 | ((uint32_t)a1 << 8) | ((uint32_t)(bits & 0xffffull) << 16), (uint32_t)(bits >> 16)};
}

}  // namespace bcnenc
}  // namespace yupvq
