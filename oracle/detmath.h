// ORACLE — TEST INFRASTRUCTURE ONLY.  Never linked into, imported by or called
// from the product library (raytracing-weekend_b200/).
//
// detmath.h — the oracle's statement of the ARITHMETIC CONTRACT for the
// transcendental functions of the path (DESIGN.md "Arithmetic contract").
//
// The reference takes sinf/cosf/logf/powf/atan2f/asinf from CUDA's libm, built
// with FastMath (RestOfLife.vcxproj:126); none of those bits can be reproduced
// here (no OptiX build, "parity unpinned"), and libm implementations differ in
// the last ulp, which the path amplifies into different random walks
// (shaders/closehit.cu:100: the shadow ray stops 5e-5 short of the light, below
// fp32 resolution at distance >= 512).  The contract therefore DEFINES each
// function as a fixed sequence of correctly rounded fp32 operations
// (+ - * / sqrtf fmaf rintf), so the launch is a pure function of its inputs.
// tests/test_detmath.py pins this file against float64 libm (<= 2 ulp for
// sin/cos/log, <= 4 ulp for atan2/asin) and against the device build bit for bit.
//
// Call sites restated with these: lib/sampling.cuh:16-19,55-56 (sin, cos);
// geometry/volumeBox.cu:79, geometry/volumeSphere.cu:93 (log);
// material/dielectricMaterial.cu:25 (powf(x,5)); geometry/sphere.cu:27-28
// (atan2, asin); texture/checkeredTexture.cu:10, texture/noiseTexture.cu:80 (sin).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace orc {
namespace dm {

static inline uint32_t bits(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static inline float from_bits(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

// Horner evaluation, one fmaf per coefficient, highest degree first.
template <int N>
static inline float horner(const float (&c)[N], float t) {
    float acc = c[0];
    for (int i = 1; i < N; i++) acc = std::fmaf(acc, t, c[i]);
    return acc;
}

// sin(r), cos(r) for |r| <= pi/4
static const float kSin[3] = {-1.9515295891e-4f, 8.3321608736e-3f, -1.6666654611e-1f};
static const float kCos[3] = {2.443315711809948e-5f, -1.388731625493765e-3f, 4.166664568298827e-2f};

static inline void sincos(float x, float* s_out, float* c_out) {
    float r;
    int quadrant;
    const float ax = std::fabs(x);
    if (ax <= 1.0e5f) {
        // Cody-Waite: x - k*(pi/2) with pi/2 = P1 + P2 + P3, one fmaf each
        const float k = std::rint(x * 0.636619772367581343f);
        r = std::fmaf(k, -1.5707962512969971e+0f, x);
        r = std::fmaf(k, -7.5497894158615964e-8f, r);
        r = std::fmaf(k, -5.3903029534742384e-15f, r);
        quadrant = (int)k & 3;
    } else if (ax <= 3.402823466e+38f) {
        // rare: fp64 reduction
        double xd = (double)x;
        if (std::fabs(xd) > 1.0e13) xd = std::fmod(xd, 6.283185307179586);  // bounded, accuracy not claimed
        const double k = std::rint(xd * 0.63661977236758134308);
        double rd = std::fma(k, -1.5707963267948966, xd);
        rd = std::fma(k, -6.123233995736766e-17, rd);
        r = (float)rd;
        quadrant = (int)(long long)(k - 4.0 * std::rint(k * 0.25)) & 3;
    } else {
        *s_out = NAN;
        *c_out = NAN;
        return;
    }
    const float z = r * r;
    const float sin_r = std::fmaf(horner(kSin, z) * z, r, r);
    const float cos_r = std::fmaf(horner(kCos, z), z * z, std::fmaf(-0.5f, z, 1.0f));
    // quadrant 0: (s, c); 1: (c, -s); 2: (-s, -c); 3: (-c, s)
    switch (quadrant) {
        case 0: *s_out = sin_r;  *c_out = cos_r;  break;
        case 1: *s_out = cos_r;  *c_out = -sin_r; break;
        case 2: *s_out = -sin_r; *c_out = -cos_r; break;
        default: *s_out = -cos_r; *c_out = sin_r; break;
    }
}
static inline float sin(float x) { float s, c; sincos(x, &s, &c); return s; }
static inline float cos(float x) { float s, c; sincos(x, &s, &c); return c; }

// log(x): x = m 2^e, m in [sqrt(1/2), sqrt(2)), polynomial in f = m - 1
static const float kLog[9] = {7.0376836292e-2f, -1.1514610310e-1f, 1.1676998740e-1f, -1.2420140846e-1f,
                              1.4249322787e-1f, -1.6668057665e-1f, 2.0000714765e-1f, -2.4999993993e-1f,
                              3.3333331174e-1f};
static inline float log(float x) {
    if (x == 0.0f) return -INFINITY;
    if (!(x > 0.0f)) return NAN;
    if (std::isinf(x)) return x;
    int e = 0;
    uint32_t ix = bits(x);
    if (ix < 0x00800000u) {  // subnormal
        x *= 8388608.0f;
        ix = bits(x);
        e = -23;
    }
    e += (int)(ix >> 23) - 126;
    float f = from_bits((ix & 0x007fffffu) | 0x3f000000u);  // mantissa in [0.5, 1)
    if (f < 0.707106781186547524f) {
        e -= 1;
        f = f + f - 1.0f;
    } else {
        f = f - 1.0f;
    }
    const float z = f * f;
    float y = horner(kLog, f) * f * z;
    const float fe = (float)e;
    y = std::fmaf(-2.12194440e-4f, fe, y);   // ln2 = 0.693359375 - 2.12194440e-4
    y = std::fmaf(-0.5f, z, y);
    return std::fmaf(0.693359375f, fe, f + y);
}

// powf(x, 5.f) of fresnelSchlick
static inline float pow5(float x) {
    const float x2 = x * x;
    return (x2 * x2) * x;
}

static const float kAtan[4] = {8.05374449538e-2f, -1.38776856032e-1f, 1.99777106478e-1f, -3.33329491539e-1f};
static inline float atan_nonneg(float t) {
    float base = 0.0f;
    if (t > 2.414213562373095f) {
        base = 1.5707963267948966f;
        t = -(1.0f / t);
    } else if (t > 0.4142135623730950f) {
        base = 0.7853981633974483f;
        t = (t - 1.0f) / (t + 1.0f);
    }
    const float z = t * t;
    return base + std::fmaf(horner(kAtan, z) * z, t, t);
}
static inline float atan2(float y, float x) {
    if (std::isnan(x) || std::isnan(y)) return NAN;
    const float ax = std::fabs(x), ay = std::fabs(y);
    float a;
    if (ax == 0.0f && ay == 0.0f) a = 0.0f;
    else if (std::isinf(ax) && std::isinf(ay)) a = 0.7853981633974483f;
    else a = atan_nonneg(ay / ax);
    if (std::signbit(x)) a = 3.14159265358979324f - a;
    return std::signbit(y) ? -a : a;
}

static const float kAsin[5] = {4.2163199048e-2f, 2.4181311049e-2f, 4.5470025998e-2f, 7.4953002686e-2f,
                               1.6666752422e-1f};
static inline float asin(float x) {
    const float a = std::fabs(x);
    if (!(a <= 1.0f)) return NAN;
    float r;
    if (a > 0.5f) {
        const float z = 0.5f * (1.0f - a);
        const float w = std::sqrt(z);
        const float h = std::fmaf(horner(kAsin, z) * z, w, w);
        r = 1.5707963267948966f - (h + h);
    } else {
        const float z = a * a;
        r = std::fmaf(horner(kAsin, z) * z, a, a);
    }
    return std::signbit(x) ? -r : r;
}

}  // namespace dm
}  // namespace orc
