// TEST INFRASTRUCTURE — CPU restatement ("port") of the reference VQ hot path.  NOT product code:
// only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// build, load or call this file.  The product (yuptexture_b200/csrc) never links or calls it.
//
// Every function cites the reference lines it restates (paths relative to /root/reference).
// Parity pin: tests/test_oracle_golden.py checks this file bit-for-bit against the golden
// vectors under tests/golden/ (generated from the unmodified reference by
// tests/golden/make_golden.py through oracle/_ref/libyupref.so), and
// tests/test_oracle_vs_ref.py checks it live against oracle/_ref when that library is present.
//
// Build: g++ -std=c++17 -O2 -fopenmp -mfma -ffp-contract=off -shared -fPIC  (oracle/Makefile).
// -ffp-contract=off + explicit std::fma pins the FMA pattern GCC chose for the reference TU
// (10 vfmadd in PixelToOklab/OklabToPixel); it must not be left to the
// optimiser.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <numeric>
#include <random>
#include <vector>

#include "../yuptexture_b200/csrc/bc7_tables.h"  // constant tables only (generated, see header)

namespace {

// ---------------------------------------------------------------------------------------------
// colour pipeline
// ---------------------------------------------------------------------------------------------

// colorm ColorBase::dot, thirdparty/colorm/include/colorm.h:1034-1037: m1*f1 + m2*f2 + m3*f3.
// GCC -O2 -mfma contracts it as fma(m3,f3, fma(m1,f1, m2*f2)) -- verified exhaustively over all
// 2^24 sRGB8 colours against oracle/_ref (tests/test_oracle_vs_ref.py::test_oklab_exhaustive).
inline double dot3(double m1, double m2, double m3, double f1, double f2, double f3) {
    return std::fma(m3, f3, std::fma(m1, f1, m2 * f2));
}

// Lrgb::linear, colorm.h:4023-4041 (value on the 0..255 scale).
inline double srgb255_to_linear(double value) {
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

// Rgb::nonLinear, colorm.h:3984-4000 (returns the 0..255 scale).
inline double linear_to_srgb255(double value) {
    if (value >= 0.0) {
        if (value <= 0.0031308) return 12.92 * value * 255.0;
        return (1.055 * std::pow(value, 1.0 / 2.4) - 0.055) * 255.0;
    }
    value = -value;
    if (value <= 0.0031308) return -(12.92 * value * 255.0);
    return -((1.055 * std::pow(value, 1.0 / 2.4) - 0.055) * 255.0);
}

// Oklab::Oklab(const Lrgb&), colorm.h:4749-4759, after Lrgb::Lrgb(const Rgb&), colorm.h:4441-4446.
inline void rgb255_to_oklab(double r, double g, double b, float* out) {
    const double lr = srgb255_to_linear(r), lg = srgb255_to_linear(g), lb = srgb255_to_linear(b);
    const double l = dot3(lr, lg, lb, 0.4122214708018041, 0.53633253634543, 0.05144599285276585);
    const double m = dot3(lr, lg, lb, 0.2119034982505858, 0.6806995451361225, 0.1073969566132915);
    const double s = dot3(lr, lg, lb, 0.08830246188874209, 0.2817188376235317, 0.6299787004877261);
    const double l2 = std::cbrt(l), m2 = std::cbrt(m), s2 = std::cbrt(s);
    out[0] = (float)dot3(l2, m2, s2, 0.2104542682745812, 0.7936177747300267, -0.004072043004608028);
    out[1] = (float)dot3(l2, m2, s2, 1.977998532388508, -2.428592241936286, 0.4505937095477779);
    out[2] = (float)dot3(l2, m2, s2, 0.02590404248765818, 0.7827717124269177, -0.8086757549145759);
}

// Rgb(Lrgb(Oklab)), colorm.h:4475-4485 then :4292-4297.  Returns the 0..255 doubles.
inline void oklab_to_rgb255(double L, double A, double B, double* rgb) {
    // dot(1.0, f2, f3) = m1*1.0 + m2*f2 + m3*f3; the *1.0 folds away, leaving two chained FMAs.
    const double l2 = std::fma(B, 0.2158037573, std::fma(A, 0.3963377774, L));
    const double m2 = std::fma(B, -0.0638541728, std::fma(A, -0.1055613458, L));
    const double s2 = std::fma(B, -1.291485548, std::fma(A, -0.0894841775, L));
    const double l = l2 * l2 * l2, m = m2 * m2 * m2, s = s2 * s2 * s2;
    const double lr = dot3(l, m, s, 4.0767416621, -3.3077115913, 0.2309699292);
    const double lg = dot3(l, m, s, -1.2684380046, 2.6097574011, -0.3413193965);
    const double lb = dot3(l, m, s, -0.0041960863, -0.7034186147, 1.707614701);
    rgb[0] = linear_to_srgb255(lr);
    rgb[1] = linear_to_srgb255(lg);
    rgb[2] = linear_to_srgb255(lb);
}

// Rgb::red8, colorm.h:1310-1313 -> double255ToInteger :263-266 -> normalize255(round(x)).
inline uint8_t to_u8(double v) {
    double r = std::round(v);
    if (!(r > 0.0)) r = 0.0;  // normalize(): clamp to [0,255]; NaN maps to 0 as the cast would
    if (r > 255.0) r = 255.0;
    return (uint8_t)(unsigned int)r;
}

// VQEncoder::PixelToOklab, src/vq_encoder.cpp:37-64.
inline void pixel_to_oklab(const uint8_t* px, int ch, float* o) {
    float alpha = 1.0f;
    double r, g, b;
    switch (ch) {
        case 1: r = g = b = px[0]; break;
        case 2: r = g = b = px[0]; alpha = px[1] / 255.0f; break;
        case 3: r = px[0]; g = px[1]; b = px[2]; break;
        case 4: r = px[0]; g = px[1]; b = px[2]; alpha = px[3] / 255.0f; break;
        default: o[0] = o[1] = o[2] = o[3] = 0; return;
    }
    rgb255_to_oklab(r, g, b, o);
    o[3] = alpha;
}

// VQEncoder::OklabToPixel, src/vq_encoder.cpp:67-93 (rgb.clip() result is discarded: no-op).
inline void oklab_to_pixel(const float* o, int ch, uint8_t* px) {
    double rgb[3];
    oklab_to_rgb255(o[0], o[1], o[2], rgb);
    const uint8_t r = to_u8(rgb[0]), g = to_u8(rgb[1]), b = to_u8(rgb[2]);
    const uint8_t a = (uint8_t)std::clamp(o[3] * 255.0f, 0.0f, 255.0f);
    // r*0.2126f + g*0.7152f + b*0.0722f, same contraction shape as dot3: fma(b,., fma(r,., g*.))
    const float luma = std::fma((float)b, 0.0722f, std::fma((float)r, 0.2126f, (float)g * 0.7152f));
    switch (ch) {
        case 1: px[0] = (uint8_t)luma; break;
        case 2: px[0] = (uint8_t)luma; px[1] = a; break;
        case 3: px[0] = r; px[1] = g; px[2] = b; break;
        case 4: px[0] = r; px[1] = g; px[2] = b; px[3] = a; break;
    }
}

// VQEncoder::RgbaFloatToOklab, src/vq_encoder.cpp:118-147 (raw floats land on colorm's 0..255 scale).
inline void float_to_oklab(const float* px, int ch, float* o) {
    float alpha = 1.0f;
    double r, g, b;
    switch (ch) {
        case 1: r = g = b = px[0]; break;
        case 2: r = g = b = px[0]; alpha = px[1]; break;
        case 3: r = px[0]; g = px[1]; b = px[2]; break;
        case 4: r = px[0]; g = px[1]; b = px[2]; alpha = px[3]; break;
        default: o[0] = o[1] = o[2] = o[3] = 0; return;
    }
    rgb255_to_oklab(r, g, b, o);
    o[3] = alpha;
}

// VQEncoder::OklabToRgbaFloat, src/vq_encoder.cpp:150-176 (no clamp).
inline void oklab_to_float(const float* o, int ch, float* px) {
    double rgb[3];
    oklab_to_rgb255(o[0], o[1], o[2], rgb);
    const float r = (float)rgb[0], g = (float)rgb[1], b = (float)rgb[2];
    const float luma = std::fma(b, 0.0722f, std::fma(r, 0.2126f, g * 0.7152f));
    switch (ch) {
        case 1: px[0] = luma; break;
        case 2: px[0] = luma; px[1] = o[3]; break;
        case 3: px[0] = r; px[1] = g; px[2] = b; break;
        case 4: px[0] = r; px[1] = g; px[2] = b; px[3] = o[3]; break;
    }
}

// ---------------------------------------------------------------------------------------------
// distances
// ---------------------------------------------------------------------------------------------

// VQEncoder::BlockDistanceSAD, src/vq_encoder.cpp:201-208.
inline float block_sad(const uint8_t* a, const uint8_t* b, int ch) {
    uint32_t sad = 0;
    for (int i = 0; i < 16 * ch; ++i) sad += (uint32_t)std::abs((int)a[i] - (int)b[i]);
    return (float)sad;
}

// VQEncoder::OklabBlockDistanceSq_SIMD, src/vq_encoder.cpp:212-237, scalar restatement of the
// AVX2 op order: 8 lane accumulators, sub -> mul by weight -> fma,
// then the fold (lane j + lane j+4) and the two hadd levels.
inline float oklab_dist(const float* a, const float* b) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = 0; i < 64; i += 8)
        for (int l = 0; l < 8; ++l) {
            const float w = (l & 3) == 0 ? 2.0f : 1.0f;
            const float d = a[i + l] - b[i + l];
            const float wd = d * w;
            acc[l] = std::fma(wd, wd, acc[l]);
        }
    const float s0 = acc[0] + acc[4], s1 = acc[1] + acc[5], s2 = acc[2] + acc[6], s3 = acc[3] + acc[7];
    return (s0 + s1) + (s2 + s3);
}

// ---------------------------------------------------------------------------------------------
// BCn block decode (Compressonator 4.3 behaviour, pinned by tests/golden)
// The reference reaches it through BCnCompressor::Decompress, src/bcn_compressor.cpp:191-227,
// called per block at src/vq_bcn_compressor.cpp:142-146.
// ---------------------------------------------------------------------------------------------

inline void rgb565(uint16_t c, int* rgb) {
    const int r = (c >> 11) & 31, g = (c >> 5) & 63, b = c & 31;
    rgb[0] = (r << 3) | (r >> 2);
    rgb[1] = (g << 2) | (g >> 4);
    rgb[2] = (b << 3) | (b >> 2);
}

// Colour half of BC1/BC3.  three_colour_allowed=false for BC3 (always 4-colour mode).
// out: 16 x RGBA (alpha 255, or 0 for the punch-through index).
void decode_bc1_colour(const uint8_t* blk, bool three_colour_allowed, uint8_t out[16][4]) {
    const uint16_t c0 = (uint16_t)(blk[0] | (blk[1] << 8)), c1 = (uint16_t)(blk[2] | (blk[3] << 8));
    int e0[3], e1[3];
    rgb565(c0, e0);
    rgb565(c1, e1);
    uint8_t pal[4][4];
    for (int k = 0; k < 3; ++k) {
        pal[0][k] = (uint8_t)e0[k];
        pal[1][k] = (uint8_t)e1[k];
    }
    pal[0][3] = pal[1][3] = pal[2][3] = pal[3][3] = 255;
    if (c0 > c1 || !three_colour_allowed) {
        for (int k = 0; k < 3; ++k) {
            pal[2][k] = (uint8_t)((2 * e0[k] + e1[k]) / 3);
            pal[3][k] = (uint8_t)((e0[k] + 2 * e1[k]) / 3);
        }
    } else {
        for (int k = 0; k < 3; ++k) {
            pal[2][k] = (uint8_t)((e0[k] + e1[k]) / 2);
            pal[3][k] = 0;
        }
        pal[3][3] = 0;
    }
    const uint32_t bits = (uint32_t)blk[4] | ((uint32_t)blk[5] << 8) | ((uint32_t)blk[6] << 16) | ((uint32_t)blk[7] << 24);
    for (int p = 0; p < 16; ++p) std::memcpy(out[p], pal[(bits >> (2 * p)) & 3], 4);
}

// Alpha half of BC3 (also the BC4 / BC5 channel block).
void decode_bc3_alpha(const uint8_t* blk, uint8_t out[16]) {
    const int a0 = blk[0], a1 = blk[1];
    int pal[8];
    pal[0] = a0;
    pal[1] = a1;
    if (a0 > a1) {
        for (int i = 2; i < 8; ++i) pal[i] = ((8 - i) * a0 + (i - 1) * a1 + 3) / 7;
    } else {
        for (int i = 2; i < 6; ++i) pal[i] = ((6 - i) * a0 + (i - 1) * a1 + 2) / 5;
        pal[6] = 0;
        pal[7] = 255;
    }
    uint64_t bits = 0;
    for (int i = 0; i < 6; ++i) bits |= (uint64_t)blk[2 + i] << (8 * i);
    for (int p = 0; p < 16; ++p) out[p] = (uint8_t)pal[(bits >> (3 * p)) & 7];
}

struct BitReader {
    const uint8_t* d;
    int pos = 0;
    explicit BitReader(const uint8_t* p) : d(p) {}
    uint32_t get(int n) {
        uint32_t v = 0;
        for (int i = 0; i < n; ++i, ++pos) v |= (uint32_t)((d[pos >> 3] >> (pos & 7)) & 1) << i;
        return v;
    }
};

const int kBc7W2[4] = {0, 21, 43, 64};
const int kBc7W3[8] = {0, 9, 18, 27, 37, 46, 55, 64};
const int kBc7W4[16] = {0, 4, 9, 13, 17, 21, 26, 30, 34, 38, 43, 47, 51, 55, 60, 64};

inline int bc7_interp(int a, int b, int w) { return (a * (64 - w) + b * w + 32) >> 6; }
inline const int* bc7_weights(int bits) { return bits == 2 ? kBc7W2 : bits == 3 ? kBc7W3 : kBc7W4; }

// BC7 (BPTC) block -> 16 RGBA pixels.
void decode_bc7(const uint8_t* blk, uint8_t out[16][4]) {
    static const int NS[8] = {3, 2, 3, 2, 1, 1, 1, 2};   // subsets
    static const int PB[8] = {4, 6, 6, 6, 0, 0, 0, 6};   // partition bits
    static const int RB[8] = {0, 0, 0, 0, 2, 2, 0, 0};   // rotation bits
    static const int ISB[8] = {0, 0, 0, 0, 1, 0, 0, 0};  // index selection bit
    static const int CB[8] = {4, 6, 5, 7, 5, 7, 7, 5};   // colour bits
    static const int AB[8] = {0, 0, 0, 0, 6, 8, 7, 5};   // alpha bits
    static const int EPB[8] = {1, 0, 0, 1, 0, 0, 1, 1};  // per-endpoint p-bit
    static const int SPB[8] = {0, 1, 0, 0, 0, 0, 0, 0};  // shared p-bit per subset
    static const int IB[8] = {3, 3, 2, 2, 2, 2, 4, 2};   // index bits
    static const int IB2[8] = {0, 0, 0, 0, 3, 2, 0, 0};  // secondary index bits

    int mode = 0;
    while (mode < 8 && !((blk[0] >> mode) & 1)) ++mode;
    if (mode >= 8) {
        std::memset(out, 0, 64);
        return;
    }
    BitReader br(blk);
    br.get(mode + 1);
    const int ns = NS[mode];
    const int part = (int)br.get(PB[mode]);
    const int rot = (int)br.get(RB[mode]);
    const int isel = (int)br.get(ISB[mode]);
    int ep[6][4];
    for (int c = 0; c < 3; ++c)
        for (int e = 0; e < 2 * ns; ++e) ep[e][c] = (int)br.get(CB[mode]);
    for (int e = 0; e < 2 * ns; ++e) ep[e][3] = AB[mode] ? (int)br.get(AB[mode]) : 255;
    int cbits = CB[mode], abits = AB[mode];
    if (EPB[mode]) {
        for (int e = 0; e < 2 * ns; ++e) {
            const int p = (int)br.get(1);
            for (int c = 0; c < 3; ++c) ep[e][c] = (ep[e][c] << 1) | p;
            if (abits) ep[e][3] = (ep[e][3] << 1) | p;
        }
        ++cbits;
        if (abits) ++abits;
    } else if (SPB[mode]) {
        for (int s = 0; s < ns; ++s) {
            const int p = (int)br.get(1);
            for (int e = 2 * s; e < 2 * s + 2; ++e)
                for (int c = 0; c < 3; ++c) ep[e][c] = (ep[e][c] << 1) | p;
        }
        ++cbits;
    }
    for (int e = 0; e < 2 * ns; ++e) {
        for (int c = 0; c < 3; ++c) {
            const int v = ep[e][c] << (8 - cbits);
            ep[e][c] = v | (v >> cbits);
        }
        if (abits) {
            const int v = ep[e][3] << (8 - abits);
            ep[e][3] = v | (v >> abits);
        }
    }
    const uint8_t* ptab = ns == 1 ? nullptr : ns == 2 ? kBc7Partition2[part] : kBc7Partition3[part];
    int anchor[3] = {0, 0, 0};
    if (ns == 2) anchor[1] = kBc7Anchor2[part];
    if (ns == 3) {
        anchor[1] = kBc7Anchor3a[part];
        anchor[2] = kBc7Anchor3b[part];
    }
    int idx1[16], idx2[16];
    for (int p = 0; p < 16; ++p) {
        const int s = ptab ? ptab[p] : 0;
        idx1[p] = (int)br.get(IB[mode] - (p == anchor[s] ? 1 : 0));
    }
    for (int p = 0; p < 16; ++p) idx2[p] = IB2[mode] ? (int)br.get(IB2[mode] - (p == 0 ? 1 : 0)) : 0;

    for (int p = 0; p < 16; ++p) {
        const int s = ptab ? ptab[p] : 0;
        const int* e0 = ep[2 * s];
        const int* e1 = ep[2 * s + 1];
        int ci = idx1[p], ai = idx1[p], cib = IB[mode], aib = IB[mode];
        if (IB2[mode]) {
            if (isel) {
                ci = idx2[p];
                cib = IB2[mode];
            } else {
                ai = idx2[p];
                aib = IB2[mode];
            }
        }
        int px[4];
        const int wc = bc7_weights(cib)[ci], wa = bc7_weights(aib)[ai];
        for (int c = 0; c < 3; ++c) px[c] = bc7_interp(e0[c], e1[c], wc);
        px[3] = bc7_interp(e0[3], e1[3], wa);
        if (rot == 1) std::swap(px[0], px[3]);
        if (rot == 2) std::swap(px[1], px[3]);
        if (rot == 3) std::swap(px[2], px[3]);
        for (int c = 0; c < 4; ++c) out[p][c] = (uint8_t)px[c];
    }
}

// Component selection of Compressonator's RGBA -> R_8 / RG_8 / RGB_888 conversion as observed
// through BCnCompressor::Decompress (src/bcn_compressor.cpp:191-227): 1-channel output takes B
// (R for BC4), 2-channel output takes (R,G) except BC5 which yields (G,B).  Quirks, replicated.
void store_pixels(const uint8_t rgba[16][4], int fmt, int ch, uint8_t* out) {
    int map[4] = {0, 1, 2, 3};
    if (ch == 1) map[0] = (fmt == 4) ? 0 : 2;
    if (ch == 2 && fmt == 5) {
        map[0] = 1;
        map[1] = 2;
    }
    for (int p = 0; p < 16; ++p)
        for (int c = 0; c < ch; ++c) out[p * ch + c] = rgba[p][map[c]];
}

// ---------------------------------------------------------------------------------------------
// k-means pieces
// ---------------------------------------------------------------------------------------------

struct Draws {
    std::vector<size_t> order;  // sampled block indices (identity when ratio >= 1)
    size_t first = 0;
    std::mt19937 rng;
};

// Sampling prologue of BuildCodebook, src/vq_encoder.cpp:262-284 (HDR :494-513).
Draws make_sample(size_t n, uint32_t K, float ratio, uint32_t seed) {
    Draws d;
    d.rng.seed(seed);
    if (ratio < 1.0f && ratio > 0.0f) {
        size_t m = (size_t)(n * ratio);
        m = std::max((size_t)K, m);
        m = std::min(m, n);
        std::vector<size_t> idx(n);
        std::iota(idx.begin(), idx.end(), (size_t)0);
        std::shuffle(idx.begin(), idx.end(), d.rng);
        d.order.assign(idx.begin(), idx.begin() + m);
    } else {
        d.order.resize(n);
        std::iota(d.order.begin(), d.order.end(), (size_t)0);
    }
    return d;
}

// argmin over centroids with strict '<' (lowest index wins), src/vq_encoder.cpp:332-337.
inline void nearest(const float* x, const float* cent, uint32_t K, uint32_t* best, float* best_d) {
    float md = std::numeric_limits<float>::max();
    uint32_t bc = 0;
    for (uint32_t c = 0; c < K; ++c) {
        const float d = oklab_dist(x, cent + (size_t)c * 64);
        if (d < md) {
            md = d;
            bc = c;
        }
    }
    *best = bc;
    *best_d = md;
}

// Lloyd iterations in Oklab space, src/vq_encoder.cpp:328-376 (LDR, min_iter_for_break = 0) and
// :559-606 (HDR, min_iter_for_break = 1).  `threads` emulates the reference's static OpenMP
// chunking of the fp32 accumulation (:345-361): thread t sums its contiguous chunk in block order
// into a private copy and the copies are merged in thread order (the reference merges in arrival
// order under `omp critical`; thread order is one of its possible outcomes, and the only one at
// numThreads = 1).  Empty clusters follow the serial semantics (ascending c, first maximum).
uint32_t lloyd(const std::vector<float>& feat, size_t n, std::vector<float>& cent, uint32_t K, uint32_t max_iter,
               uint32_t break_after, int threads, std::vector<uint32_t>& assign) {
    std::vector<float> errors(n);
    assign.assign(n, 0);
    uint32_t iters = 0;
    if (threads < 1) threads = 1;
    for (uint32_t iter = 0; iter < max_iter; ++iter) {
        bool changed = false;
#pragma omp parallel for reduction(|| : changed) schedule(static)
        for (int64_t i = 0; i < (int64_t)n; ++i) {
            uint32_t bc;
            float md;
            nearest(&feat[(size_t)i * 64], cent.data(), K, &bc, &md);
            errors[i] = md;
            if (assign[i] != bc) {
                assign[i] = bc;
                changed = true;
            }
        }
        ++iters;
        if (!changed && iter > break_after) break;  // LDR: iter > 0 (:341); HDR: iter > 1 (:572)

        std::vector<float> sums((size_t)K * 64, 0.0f);
        std::vector<uint32_t> counts(K, 0);
        const size_t q = n / (size_t)threads, r = n % (size_t)threads;
        size_t begin = 0;
        for (int t = 0; t < threads; ++t) {
            const size_t len = q + ((size_t)t < r ? 1 : 0);
            std::vector<float> ls((size_t)K * 64, 0.0f);
            std::vector<uint32_t> lc(K, 0);
            for (size_t i = begin; i < begin + len; ++i) {
                const uint32_t c = assign[i];
                lc[c]++;
                for (int j = 0; j < 64; ++j) ls[(size_t)c * 64 + j] += feat[i * 64 + j];
            }
            for (uint32_t c = 0; c < K; ++c) {
                counts[c] += lc[c];
                for (int j = 0; j < 64; ++j) sums[(size_t)c * 64 + j] += ls[(size_t)c * 64 + j];
            }
            begin += len;
        }
        for (uint32_t c = 0; c < K; ++c) {
            if (counts[c] > 0) {
                const float inv = 1.0f / counts[c];
                for (int j = 0; j < 64; ++j) cent[(size_t)c * 64 + j] = sums[(size_t)c * 64 + j] * inv;
            } else {
                const size_t worst = (size_t)(std::max_element(errors.begin(), errors.end()) - errors.begin());
                std::memcpy(&cent[(size_t)c * 64], &feat[worst * 64], 64 * sizeof(float));
                errors[worst] = 0.0f;
            }
        }
    }
    return iters;
}

}  // namespace

extern "C" {

// CompressionConfig::SetQuality, src/vq_encoder.cpp:12-18 with the hard-wired powers 4..10
// (the params' min/max powers are copied after SetQuality and never used).
uint32_t orc_codebook_size(float quality) {
    const float q = std::clamp(quality, 0.0f, 1.0f);
    const uint32_t power = 4u + (uint32_t)roundf(q * (10 - 4));
    return 1u << power;
}

// BCn blocks -> per-block pixel layout (n x 16*ch bytes).  fmt: 1 BC1, 3 BC3, 4 BC4, 5 BC5, 7 BC7.
int orc_decode_bcn(const uint8_t* bc, size_t n, int fmt, int ch, uint8_t* out) {
    if (ch < 1 || ch > 4) return 1;
#pragma omp parallel for
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        uint8_t rgba[16][4];
        switch (fmt) {
            case 1:
                decode_bc1_colour(bc + i * 8, true, rgba);  // ch==3 drops alpha: punch-through -> black
                break;
            case 3: {
                decode_bc1_colour(bc + i * 16 + 8, false, rgba);
                uint8_t a[16];
                decode_bc3_alpha(bc + i * 16, a);
                for (int p = 0; p < 16; ++p) rgba[p][3] = a[p];
                break;
            }
            case 4: {
                uint8_t a[16];
                decode_bc3_alpha(bc + i * 8, a);
                for (int p = 0; p < 16; ++p) {
                    rgba[p][0] = rgba[p][1] = rgba[p][2] = a[p];
                    rgba[p][3] = 255;
                }
                break;
            }
            case 5: {
                uint8_t a[16], b[16];
                decode_bc3_alpha(bc + i * 16, a);
                decode_bc3_alpha(bc + i * 16 + 8, b);
                for (int p = 0; p < 16; ++p) {
                    rgba[p][0] = a[p];
                    rgba[p][1] = b[p];
                    rgba[p][2] = 0;
                    rgba[p][3] = 255;
                }
                break;
            }
            case 7:
                decode_bc7(bc + i * 16, rgba);
                break;
            default:
                std::memset(rgba, 0, sizeof(rgba));
        }
        store_pixels(rgba, fmt, ch, out + (size_t)i * 16 * ch);
    }
    return (fmt == 1 || fmt == 3 || fmt == 4 || fmt == 5 || fmt == 7) ? 0 : 2;
}

// HDR tile gather, src/vq_bcn_compressor.cpp:235-252.
void orc_gather_tiles_f32(const float* img, uint32_t w, uint32_t h, uint32_t ch, float* out) {
    const size_t bx = (w + 3) / 4, by = (h + 3) / 4;
    for (size_t b = 0; b < bx * by; ++b) {
        const size_t x0 = (b % bx) * 4, y0 = (b / bx) * 4;
        for (size_t y = 0; y < 4; ++y)
            for (size_t x = 0; x < 4; ++x) {
                const size_t px = std::min(x0 + x, (size_t)w - 1), py = std::min(y0 + y, (size_t)h - 1);
                std::memcpy(out + (b * 16 + y * 4 + x) * ch, img + (py * w + px) * ch, ch * sizeof(float));
            }
    }
}

// PixelBlockToOklabBlock, src/vq_encoder.cpp:96-102.
void orc_pixels_to_oklab(const uint8_t* px, size_t n, int ch, float* out) {
#pragma omp parallel for
    for (int64_t i = 0; i < (int64_t)n; ++i)
        for (int p = 0; p < 16; ++p) pixel_to_oklab(px + ((size_t)i * 16 + p) * ch, ch, out + (size_t)i * 64 + p * 4);
}

// OklabBlockToPixelBlock, src/vq_encoder.cpp:105-111.
void orc_oklab_to_pixels(const float* ok, size_t n, int ch, uint8_t* out) {
    for (size_t i = 0; i < n; ++i)
        for (int p = 0; p < 16; ++p) oklab_to_pixel(ok + i * 64 + p * 4, ch, out + (i * 16 + p) * ch);
}

// RgbaFloatBlockToOklabBlock, src/vq_encoder.cpp:179-185.
void orc_floats_to_oklab(const float* px, size_t n, int ch, float* out) {
#pragma omp parallel for
    for (int64_t i = 0; i < (int64_t)n; ++i)
        for (int p = 0; p < 16; ++p) float_to_oklab(px + ((size_t)i * 16 + p) * ch, ch, out + (size_t)i * 64 + p * 4);
}

// OklabBlockToRgbaFloatBlock, src/vq_encoder.cpp:188-194.
void orc_oklab_to_floats(const float* ok, size_t n, int ch, float* out) {
    for (size_t i = 0; i < n; ++i)
        for (int p = 0; p < 16; ++p) oklab_to_float(ok + i * 64 + p * 4, ch, out + (i * 16 + p) * ch);
}

void orc_oklab_distance(const float* a, size_t na, const float* b, size_t nb, float* out) {
#pragma omp parallel for
    for (int64_t i = 0; i < (int64_t)na; ++i)
        for (size_t j = 0; j < nb; ++j) out[(size_t)i * nb + j] = oklab_dist(a + (size_t)i * 64, b + j * 64);
}

void orc_block_sad(const uint8_t* a, size_t na, const uint8_t* b, size_t nb, int ch, float* out) {
    for (size_t i = 0; i < na; ++i)
        for (size_t j = 0; j < nb; ++j) out[i * nb + j] = block_sad(a + i * 16 * ch, b + j * 16 * ch, ch);
}

// Exact assignment of Oklab feature vectors (the loop of src/vq_encoder.cpp:330-340 / :462-471).
void orc_assign(const float* feat, size_t n, const float* cent, uint32_t K, uint32_t* idx, float* err) {
#pragma omp parallel for
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        uint32_t bc;
        float md;
        nearest(feat + (size_t)i * 64, cent, K, &bc, &md);
        idx[i] = bc;
        if (err) err[i] = md;
    }
}

// K-Means++ seeding on uint8 pixel blocks with SAD^2, src/vq_encoder.cpp:287-313.
// out_seed_idx[k] = index (into the sampled order) of the block chosen as seed k.
// Returns the number of rounds actually drawn (K-1 unless the sum hit zero).
uint32_t orc_seed_ldr(const uint8_t* px, size_t n, int ch, uint32_t K, uint32_t seed, uint32_t* out_seed_idx) {
    std::mt19937 rng(seed);
    std::vector<float> minD(n, std::numeric_limits<float>::max());
    std::uniform_int_distribution<size_t> distrib(0, n - 1);
    out_seed_idx[0] = (uint32_t)distrib(rng);
    uint32_t rounds = 0;
    for (uint32_t i = 1; i < K; ++i) {
        double sum = 0.0;
        const uint8_t* prev = px + (size_t)out_seed_idx[i - 1] * 16 * ch;
#pragma omp parallel for reduction(+ : sum)
        for (int64_t j = 0; j < (int64_t)n; ++j) {
            const float d = block_sad(px + (size_t)j * 16 * ch, prev, ch);
            minD[j] = std::min(d * d, minD[j]);
            sum += minD[j];
        }
        if (sum <= 0) {
            for (uint32_t k = i; k < K; ++k) out_seed_idx[k] = out_seed_idx[0];
            break;
        }
        ++rounds;
        std::uniform_real_distribution<double> pd(0.0, sum);
        const double p = pd(rng);
        double cum = 0.0;
        // The reference leaves the seed empty if the scan falls through (cannot happen: cum == sum >= p).
        out_seed_idx[i] = (uint32_t)(n - 1);
        for (size_t j = 0; j < n; ++j) {
            cum += minD[j];
            if (cum >= p) {
                out_seed_idx[i] = (uint32_t)j;
                break;
            }
        }
    }
    return rounds;
}

// BuildCodebook, src/vq_encoder.cpp:260-441 up to (not including) the Compressonator centroid
// encode (A16 stays on the host in both the reference and the product).
// metric: 0 SAD, 1 Oklab.  out_centroids: K x 16*ch bytes.  out_iters: Lloyd assign passes run.
int orc_build_codebook(const uint8_t* px_all, size_t n_all, int ch, uint32_t K, uint32_t max_iter, int metric,
                       float ratio, uint32_t seed, int threads, uint8_t* out_centroids, uint32_t* out_iters) {
    if (n_all == 0 || ch < 1 || ch > 4) return 1;
    const size_t bs = 16 * (size_t)ch;
    Draws dr = make_sample(n_all, K, ratio, seed);
    const size_t n = dr.order.size();
    std::vector<uint8_t> px(n * bs);
    for (size_t i = 0; i < n; ++i) std::memcpy(&px[i * bs], px_all + dr.order[i] * bs, bs);

    // --- K-Means++ (always SAD), :287-313
    std::vector<uint8_t> pc((size_t)K * bs);
    std::vector<float> minD(n, std::numeric_limits<float>::max());
    std::uniform_int_distribution<size_t> distrib(0, n - 1);
    std::memcpy(&pc[0], &px[distrib(dr.rng) * bs], bs);
    for (uint32_t i = 1; i < K; ++i) {
        double sum = 0.0;
        const uint8_t* prev = &pc[(size_t)(i - 1) * bs];
#pragma omp parallel for reduction(+ : sum)
        for (int64_t j = 0; j < (int64_t)n; ++j) {
            const float d = block_sad(&px[(size_t)j * bs], prev, ch);
            minD[j] = std::min(d * d, minD[j]);
            sum += minD[j];
        }
        if (sum <= 0) {
            for (uint32_t k = i; k < K; ++k) std::memcpy(&pc[(size_t)k * bs], &pc[0], bs);
            break;
        }
        std::uniform_real_distribution<double> pd(0.0, sum);
        const double p = pd(dr.rng);
        double cum = 0.0;
        for (size_t j = 0; j < n; ++j) {
            cum += minD[j];
            if (cum >= p) {
                std::memcpy(&pc[(size_t)i * bs], &px[j * bs], bs);
                break;
            }
        }
    }

    uint32_t iters = 0;
    if (metric == 1) {
        std::vector<float> feat(n * 64), cent((size_t)K * 64);
        orc_pixels_to_oklab(px.data(), n, ch, feat.data());
        orc_pixels_to_oklab(pc.data(), K, ch, cent.data());
        std::vector<uint32_t> assign;
        iters = lloyd(feat, n, cent, K, max_iter, 0, threads, assign);
        orc_oklab_to_pixels(cent.data(), K, ch, pc.data());  // :377-378
    } else {
        // SAD-metric Lloyd, src/vq_encoder.cpp:380-430.
        std::vector<uint32_t> assign(n, 0);
        std::vector<float> errors(n);
        for (uint32_t iter = 0; iter < max_iter; ++iter) {
            bool changed = false;
#pragma omp parallel for reduction(|| : changed)
            for (int64_t i = 0; i < (int64_t)n; ++i) {
                float md = std::numeric_limits<float>::max();
                uint32_t bc = 0;
                for (uint32_t c = 0; c < K; ++c) {
                    const float d = block_sad(&px[(size_t)i * bs], &pc[(size_t)c * bs], ch);
                    if (d < md) {
                        md = d;
                        bc = c;
                    }
                }
                errors[i] = md;
                if (assign[i] != bc) {
                    assign[i] = bc;
                    changed = true;
                }
            }
            ++iters;
            if (!changed && iter > 0) break;
            std::vector<uint64_t> sums((size_t)K * bs, 0);
            std::vector<uint32_t> counts(K, 0);
            for (size_t i = 0; i < n; ++i) {
                counts[assign[i]]++;
                for (size_t j = 0; j < bs; ++j) sums[(size_t)assign[i] * bs + j] += px[i * bs + j];
            }
            for (uint32_t c = 0; c < K; ++c) {
                if (counts[c] > 0) {
                    for (size_t j = 0; j < bs; ++j) pc[(size_t)c * bs + j] = (uint8_t)(sums[(size_t)c * bs + j] / counts[c]);
                } else {
                    const size_t worst = (size_t)(std::max_element(errors.begin(), errors.end()) - errors.begin());
                    std::memcpy(&pc[(size_t)c * bs], &px[worst * bs], bs);
                    errors[worst] = 0.0f;
                }
            }
        }
    }
    std::memcpy(out_centroids, pc.data(), pc.size());
    if (out_iters) *out_iters = iters;
    return 0;
}

// QuantizeBlocks, src/vq_encoder.cpp:447-486.
int orc_quantize_blocks(const uint8_t* px, size_t n, int ch, const uint8_t* cent_px, uint32_t K, int metric,
                        uint32_t* out_idx) {
    if (n == 0) return 0;
    const size_t bs = 16 * (size_t)ch;
    if (metric == 1) {
        std::vector<float> feat(n * 64), cent((size_t)K * 64);
        orc_pixels_to_oklab(px, n, ch, feat.data());
        orc_pixels_to_oklab(cent_px, K, ch, cent.data());
        orc_assign(feat.data(), n, cent.data(), K, out_idx, nullptr);
    } else {
#pragma omp parallel for
        for (int64_t i = 0; i < (int64_t)n; ++i) {
            float md = std::numeric_limits<float>::max();
            uint32_t bc = 0;
            for (uint32_t c = 0; c < K; ++c) {
                const float d = block_sad(px + (size_t)i * bs, cent_px + (size_t)c * bs, ch);
                if (d < md) {
                    md = d;
                    bc = c;
                }
            }
            out_idx[i] = bc;
        }
    }
    return 0;
}

// HDR seeding only (src/vq_encoder.cpp:526-554), exposed for kernel-level parity tests.
uint32_t orc_seed_hdr(const float* feat, size_t n, uint32_t K, uint32_t seed, uint32_t* out_seed_idx) {
    std::mt19937 rng(seed);
    std::vector<float> minD(n, std::numeric_limits<float>::max());
    std::uniform_int_distribution<size_t> distrib(0, n - 1);
    out_seed_idx[0] = (uint32_t)distrib(rng);
    uint32_t rounds = 0;
    for (uint32_t i = 1; i < K; ++i) {
        double sum = 0.0;
        const float* prev = feat + (size_t)out_seed_idx[i - 1] * 64;
        // serial sum: the reference's OpenMP reduction order is thread-count dependent; the serial
        // order is its numThreads = 1 behaviour.
        for (size_t j = 0; j < n; ++j) {
            const float d = oklab_dist(feat + j * 64, prev);
            minD[j] = std::min(d, minD[j]);
            sum += minD[j];
        }
        if (sum <= 0) {
            for (uint32_t k = i; k < K; ++k) out_seed_idx[k] = out_seed_idx[0];
            break;
        }
        ++rounds;
        std::uniform_real_distribution<double> pd(0.0, sum);
        const double p = pd(rng);
        double cum = 0.0;
        out_seed_idx[i] = (uint32_t)(n - 1);
        for (size_t j = 0; j < n; ++j) {
            cum += minD[j];
            if (cum >= p) {
                out_seed_idx[i] = (uint32_t)j;
                break;
            }
        }
    }
    return rounds;
}

// BuildCodebookHDR, src/vq_encoder.cpp:493-622 up to the BC6H centroid encode.
int orc_build_codebook_hdr(const float* px_all, size_t n_all, int ch, uint32_t K, uint32_t max_iter, float ratio,
                           uint32_t seed, int threads, float* out_centroids, uint32_t* out_iters) {
    if (n_all == 0) return 0;  // :517
    const size_t bs = 16 * (size_t)ch;
    Draws dr = make_sample(n_all, K, ratio, seed);
    const size_t n = dr.order.size();
    std::vector<float> px(n * bs);
    for (size_t i = 0; i < n; ++i) std::memcpy(&px[i * bs], px_all + dr.order[i] * bs, bs * sizeof(float));
    std::vector<float> feat(n * 64), cent((size_t)K * 64);
    orc_floats_to_oklab(px.data(), n, ch, feat.data());

    std::vector<float> minD(n, std::numeric_limits<float>::max());
    std::uniform_int_distribution<size_t> distrib(0, n - 1);
    std::memcpy(&cent[0], &feat[distrib(dr.rng) * 64], 64 * sizeof(float));
    for (uint32_t i = 1; i < K; ++i) {
        double sum = 0.0;
        const float* prev = &cent[(size_t)(i - 1) * 64];
        for (size_t j = 0; j < n; ++j) {
            const float d = oklab_dist(&feat[j * 64], prev);
            minD[j] = std::min(d, minD[j]);
            sum += minD[j];
        }
        if (sum <= 0) {
            for (uint32_t k = i; k < K; ++k) std::memcpy(&cent[(size_t)k * 64], &cent[0], 64 * sizeof(float));
            break;
        }
        std::uniform_real_distribution<double> pd(0.0, sum);
        const double p = pd(dr.rng);
        double cum = 0.0;
        for (size_t j = 0; j < n; ++j) {
            cum += minD[j];
            if (cum >= p) {
                std::memcpy(&cent[(size_t)i * 64], &feat[j * 64], 64 * sizeof(float));
                break;
            }
        }
    }
    std::vector<uint32_t> assign;
    const uint32_t iters = lloyd(feat, n, cent, K, max_iter, 1, threads, assign);
    orc_oklab_to_floats(cent.data(), K, ch, out_centroids);  // :609-613
    if (out_iters) *out_iters = iters;
    return 0;
}

// QuantizeBlocksHDR, src/vq_encoder.cpp:629-658.
int orc_quantize_blocks_hdr(const float* px, size_t n, int ch, const float* cent_px, uint32_t K, uint32_t* out_idx) {
    if (n == 0) return 0;
    std::vector<float> feat(n * 64), cent((size_t)K * 64);
    orc_floats_to_oklab(px, n, ch, feat.data());
    orc_floats_to_oklab(cent_px, K, ch, cent.data());
    orc_assign(feat.data(), n, cent.data(), K, out_idx, nullptr);
    return 0;
}

}  // extern "C"
