// Device restatement of the slice of colorm the reference uses (all double, cast to float on
// store): reference thirdparty/colorm/include/colorm.h:4023-4041 (Lrgb::linear), :4749-4759
// (Oklab(Lrgb)), :4475-4485 (Lrgb(Oklab)), :3984-4000 (Rgb::nonLinear), :263-266 / :1310-1313
// (red8), as called by reference src/vq_encoder.cpp:37-64, :67-93, :118-147, :150-176.
//
// Bit-exactness notes:
//  * dot(): GCC -O2 -mfma contracts m1*f1 + m2*f2 + m3*f3 to fma(m3,f3, fma(m1,f1, m2*f2)) in
//    the reference TU; the same shape is spelled out here with __fma_rn/__dmul_rn.
//  * cbrt: glibc 2.39's algorithm (polynomial seed + one Halley step, no FMA) restated op for op;
//    it is not correctly rounded, so CUDA's cbrt() would differ in the last ulp.
//  * sRGB8 -> linear: 256-entry table computed on the host with the host's libm pow (the same
//    function the reference calls) and uploaded to __constant__ memory.
//  * float -> linear and linear -> sRGB need pow() on arbitrary arguments: CUDA's pow (<= 2 ulp in
//    double) is used; after the cast to float / the round to uint8 a difference is a ~2^-29 event.
#pragma once
#include "common.cuh"

namespace yupvq {

// Defined once: colour.cuh is included by convert.cu only (no relocatable device code needed).
__constant__ double c_srgb8_to_linear[256];

__device__ __forceinline__ double dot3(double m1, double m2, double m3, double f1, double f2, double f3) {
    return __fma_rn(m3, f3, __fma_rn(m1, f1, __dmul_rn(m2, f2)));
}

__device__ __forceinline__ double glibc_cbrt(double x) {
    const double ax = fabs(x);
    int xe;
    double xm;
    const unsigned long long bits = (unsigned long long)__double_as_longlong(ax);
    const int be = (int)(bits >> 52);
    if (be > 0 && be < 0x7ff) {  // normal: frexp by bit surgery
        xe = be - 1022;
        xm = __longlong_as_double((long long)((bits & 0x000fffffffffffffULL) | 0x3fe0000000000000ULL));
    } else {
        if (ax == 0.0 || be == 0x7ff) return __dadd_rn(x, x);  // 0, inf, nan
        xm = frexp(ax, &xe);                                   // subnormal
    }
    double u = __dsub_rn(0.784932344976639262, __dmul_rn(0.145263899385486377, xm));
    u = __dsub_rn(1.83469277483613086, __dmul_rn(u, xm));
    u = __dsub_rn(2.44693122563534430, __dmul_rn(u, xm));
    u = __dsub_rn(2.11499494167371287, __dmul_rn(u, xm));
    u = __dsub_rn(1.50819193781584896, __dmul_rn(u, xm));
    u = __dadd_rn(0.354895765043919860, __dmul_rn(u, xm));
    const double t2 = __dmul_rn(__dmul_rn(u, u), u);
    const int r = xe % 3;  // C remainder: sign follows xe
    const double f = r == -2 ? 1.0 / 1.5874010519681994748
                   : r == -1 ? 1.0 / 1.2599210498948731648
                   : r == 0  ? 1.0
                   : r == 1  ? 1.2599210498948731648
                             : 1.5874010519681994748;
    double ym = __dmul_rn(u, __dadd_rn(t2, __dmul_rn(2.0, xm)));
    ym = __ddiv_rn(ym, __dadd_rn(__dmul_rn(2.0, t2), xm));
    ym = __dmul_rn(ym, f);
    if (!(x > 0.0)) ym = -ym;
    return ldexp(ym, xe / 3);
}

// Lrgb::linear for an arbitrary value on the 0..255 scale (HDR path).
__device__ __forceinline__ double srgb255_to_linear(double value) {
    if (value > 0.0) {
        value = __ddiv_rn(value, 255.0);
        if (value <= 0.04045) return __ddiv_rn(value, 12.92);
        return pow(__ddiv_rn(__dadd_rn(value, 0.055), 1.055), 2.4);
    }
    value = -value;
    value = __ddiv_rn(value, 255.0);
    if (value <= 0.04045) return -__ddiv_rn(value, 12.92);
    return -pow(__ddiv_rn(__dadd_rn(value, 0.055), 1.055), 2.4);
}

// Rgb::nonLinear -> 0..255 scale.
__device__ __forceinline__ double linear_to_srgb255(double value) {
    double s = 1.0;
    if (!(value >= 0.0)) {
        value = -value;
        s = -1.0;
    }
    double r;
    if (value <= 0.0031308) r = __dmul_rn(__dmul_rn(12.92, value), 255.0);
    else r = __dmul_rn(__dsub_rn(__dmul_rn(1.055, pow(value, 1.0 / 2.4)), 0.055), 255.0);
    return s * r;
}

// linear RGB -> Oklab (L, a, b) as floats.
__device__ __forceinline__ float3 linear_to_oklab(double lr, double lg, double lb) {
    const double l = dot3(lr, lg, lb, 0.4122214708018041, 0.53633253634543, 0.05144599285276585);
    const double m = dot3(lr, lg, lb, 0.2119034982505858, 0.6806995451361225, 0.1073969566132915);
    const double s = dot3(lr, lg, lb, 0.08830246188874209, 0.2817188376235317, 0.6299787004877261);
    const double l2 = glibc_cbrt(l), m2 = glibc_cbrt(m), s2 = glibc_cbrt(s);
    float3 o;
    o.x = (float)dot3(l2, m2, s2, 0.2104542682745812, 0.7936177747300267, -0.004072043004608028);
    o.y = (float)dot3(l2, m2, s2, 1.977998532388508, -2.428592241936286, 0.4505937095477779);
    o.z = (float)dot3(l2, m2, s2, 0.02590404248765818, 0.7827717124269177, -0.8086757549145759);
    return o;
}

// sRGB8 -> Oklab, VQEncoder::PixelToOklab for the colour part.
__device__ __forceinline__ float3 srgb8_to_oklab(unsigned r, unsigned g, unsigned b) {
    return linear_to_oklab(c_srgb8_to_linear[r], c_srgb8_to_linear[g], c_srgb8_to_linear[b]);
}

// raw floats on colorm's 0..255 scale -> Oklab, VQEncoder::RgbaFloatToOklab for the colour part.
__device__ __forceinline__ float3 float_to_oklab(float r, float g, float b) {
    return linear_to_oklab(srgb255_to_linear((double)r), srgb255_to_linear((double)g), srgb255_to_linear((double)b));
}

// Oklab -> sRGB on the 0..255 scale (doubles), Rgb(Lrgb(Oklab)).
__device__ __forceinline__ void oklab_to_rgb255(float L, float A, float B, double* rgb) {
    const double l2 = __fma_rn((double)B, 0.2158037573, __fma_rn((double)A, 0.3963377774, (double)L));
    const double m2 = __fma_rn((double)B, -0.0638541728, __fma_rn((double)A, -0.1055613458, (double)L));
    const double s2 = __fma_rn((double)B, -1.291485548, __fma_rn((double)A, -0.0894841775, (double)L));
    const double l = __dmul_rn(__dmul_rn(l2, l2), l2), m = __dmul_rn(__dmul_rn(m2, m2), m2),
                 s = __dmul_rn(__dmul_rn(s2, s2), s2);
    rgb[0] = linear_to_srgb255(dot3(l, m, s, 4.0767416621, -3.3077115913, 0.2309699292));
    rgb[1] = linear_to_srgb255(dot3(l, m, s, -1.2684380046, 2.6097574011, -0.3413193965));
    rgb[2] = linear_to_srgb255(dot3(l, m, s, -0.0041960863, -0.7034186147, 1.707614701));
}

// Rgb::red8(): clamp(round(x), 0, 255), round half away from zero.
__device__ __forceinline__ unsigned to_u8(double v) {
    double r = round(v);
    if (!(r > 0.0)) r = 0.0;
    if (r > 255.0) r = 255.0;
    return (unsigned)r;
}

// r*0.2126f + g*0.7152f + b*0.0722f with the reference TU's contraction shape.
__device__ __forceinline__ float luma709(float r, float g, float b) {
    return __fmaf_rn(b, 0.0722f, __fmaf_rn(r, 0.2126f, __fmul_rn(g, 0.7152f)));
}

}  // namespace yupvq
