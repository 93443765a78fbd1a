// detmath.h — the transcendental functions of the path, defined by arithmetic.
//
// The reference calls sinf / cosf / logf / powf(x,5) / atan2f / asinf from CUDA's
// libm (lib/sampling.cuh:16-19,55-56; geometry/volumeBox.cu:79; volumeSphere.cu:93;
// material/dielectricMaterial.cu:25; geometry/sphere.cu:27-28; texture/*.cu).
// Their last bits differ between libm implementations (CUDA, glibc, fast-math),
// and the path is chaotic in them: the NEE shadow ray ends 5e-5 before the light
// (shaders/closehit.cu:100), less than one fp32 ulp at distances >= 512, so a
// one-ulp change of a bounce direction flips occlusion for ~1 % of paths.
//
// To make the whole launch a *function* of its inputs — identical bits on the
// GPU and in the CPU oracle — these functions are fixed here as explicit fp32
// sequences of + - * / sqrtf fmaf rintf (each correctly rounded on both sides).
// Accuracy is <= 2 ulp on the ranges the path uses (tests/test_detmath.py checks
// against float64), i.e. the same class as CUDA libm.  DESIGN.md "Arithmetic contract".
#pragma once
#include "vec.h"

namespace rtw {

// ---- sin / cos -------------------------------------------------------------
// Cody-Waite reduction by pi/2 in three fmaf steps for |x| <= 1e5 (every argument
// the five scenes produce), a two-step fp64 reduction beyond that (rare, exact to
// |x| ~ 1e13; larger arguments are first folded by an exact fmod), then minimax polynomials on [-pi/4, pi/4].  Non-finite x → NaN.
RTW_HD float det_sin_poly(float r) {
    const float z = r * r;
    float p = fmaf(-1.9515295891e-4f, z, 8.3321608736e-3f);
    p = fmaf(p, z, -1.6666654611e-1f);
    return fmaf(p * z, r, r);
}
RTW_HD float det_cos_poly(float r) {
    const float z = r * r;
    float p = fmaf(2.443315711809948e-5f, z, -1.388731625493765e-3f);
    p = fmaf(p, z, 4.166664568298827e-2f);
    return fmaf(p, z * z, fmaf(-0.5f, z, 1.0f));
}
RTW_HD void det_sincosf(float x, float& s, float& c) {
    float r;
    int q;
    if (fabsf(x) <= 1.0e5f) {
        const float kf = rintf(x * 0.636619772367581343f);      // x * 2/pi
        r = fmaf(kf, -1.5707962512969971e+0f, x);                // pi/2 split hi/mid/lo
        r = fmaf(kf, -7.5497894158615964e-8f, r);
        r = fmaf(kf, -5.3903029534742384e-15f, r);
        q = (int)kf & 3;
    } else {
        if (!(fabsf(x) <= RTW_FLT_MAX)) { s = NAN; c = NAN; return; }
        double xd = (double)x;
        if (fabs(xd) > 1.0e13) xd = fmod(xd, 6.283185307179586);   // exact op; accuracy not claimed here, only |result| <= 1
        const double kd = rint(xd * 0.63661977236758134308);
        double rd = fma(kd, -1.5707963267948966, xd);             // pi/2 hi (53 bits)
        rd = fma(kd, -6.123233995736766e-17, rd);                 // pi/2 lo
        r = (float)rd;
        q = (int)(long long)(kd - 4.0 * rint(kd * 0.25)) & 3;
    }
    const float sp = det_sin_poly(r), cp = det_cos_poly(r);
    const float s0 = (q & 1) ? cp : sp;
    const float c0 = (q & 1) ? sp : cp;
    s = (q & 2) ? -s0 : s0;
    c = ((q + 1) & 2) ? -c0 : c0;
}
RTW_HD float det_sinf(float x) { float s, c; det_sincosf(x, s, c); return s; }
RTW_HD float det_cosf(float x) { float s, c; det_sincosf(x, s, c); return c; }

// ---- log ---------------------------------------------------------------------
// x = m * 2^e with m in [sqrt(1/2), sqrt(2)); degree-9 polynomial in m - 1.
// log(0) = -inf, log(x < 0) = NaN, log(inf) = inf, like libm.
RTW_HD float det_logf(float x) {
    if (!(x > 0.0f)) return (x == 0.0f) ? -INFINITY : NAN;
    uint32_t ix = f2u(x);
    if (ix >= 0x7f800000u) return x;                             // +inf
    int e = 0;
    if (ix < 0x00800000u) {                                      // subnormal: scale by 2^23
        x = x * 8388608.0f;
        ix = f2u(x);
        e = -23;
    }
    e += (int)(ix >> 23) - 126;
    float m = u2f((ix & 0x007fffffu) | 0x3f000000u);             // [0.5, 1)
    if (m < 0.707106781186547524f) {
        e -= 1;
        m = m + m - 1.0f;
    } else {
        m = m - 1.0f;
    }
    const float z = m * m;
    float p = fmaf(7.0376836292e-2f, m, -1.1514610310e-1f);
    p = fmaf(p, m, 1.1676998740e-1f);
    p = fmaf(p, m, -1.2420140846e-1f);
    p = fmaf(p, m, 1.4249322787e-1f);
    p = fmaf(p, m, -1.6668057665e-1f);
    p = fmaf(p, m, 2.0000714765e-1f);
    p = fmaf(p, m, -2.4999993993e-1f);
    p = fmaf(p, m, 3.3333331174e-1f);
    float y = p * m * z;
    const float fe = (float)e;
    y = fmaf(-2.12194440e-4f, fe, y);
    y = fmaf(-0.5f, z, y);
    float r = m + y;
    r = fmaf(0.693359375f, fe, r);
    return r;
}

// ---- x^5 (fresnelSchlick's powf(x, 5.f), material/dielectricMaterial.cu:25) ----
RTW_HD float det_pow5f(float x) {
    const float x2 = x * x;
    return (x2 * x2) * x;
}

// ---- atan / atan2 / asin (get_sphere_uv, geometry/sphere.cu:24-32) -------------
RTW_HD float det_atan_pos(float x) {                             // x >= 0
    float y0;
    if (x > 2.414213562373095f) {                                // tan(3pi/8)
        y0 = 1.5707963267948966f;
        x = -(1.0f / x);
    } else if (x > 0.4142135623730950f) {                        // tan(pi/8)
        y0 = 0.7853981633974483f;
        x = (x - 1.0f) / (x + 1.0f);
    } else {
        y0 = 0.0f;
    }
    const float z = x * x;
    float p = fmaf(8.05374449538e-2f, z, -1.38776856032e-1f);
    p = fmaf(p, z, 1.99777106478e-1f);
    p = fmaf(p, z, -3.33329491539e-1f);
    return y0 + fmaf(p * z, x, x);
}
RTW_HD float det_atan2f(float y, float x) {
    if (x != x || y != y) return NAN;
    const float ax = fabsf(x), ay = fabsf(y);
    float a;
    if (ax == 0.0f && ay == 0.0f) a = 0.0f;
    else if (ax == INFINITY && ay == INFINITY) a = 0.7853981633974483f;
    else a = det_atan_pos(ay / ax);                              // ay/0 = inf → pi/2
    if (f2u(x) >> 31) a = 3.14159265358979324f - a;              // x negative (or -0)
    return (f2u(y) >> 31) ? -a : a;
}
RTW_HD float det_asinf(float x) {
    const float a = fabsf(x);
    if (!(a <= 1.0f)) return NAN;                                // like libm: |x| > 1 → NaN
    float z, w;
    const bool big = a > 0.5f;
    if (big) {
        z = 0.5f * (1.0f - a);
        w = sqrtf(z);
    } else {
        w = a;
        z = w * w;
    }
    float p = fmaf(4.2163199048e-2f, z, 2.4181311049e-2f);
    p = fmaf(p, z, 4.5470025998e-2f);
    p = fmaf(p, z, 7.4953002686e-2f);
    p = fmaf(p, z, 1.6666752422e-1f);
    float r = fmaf(p * z, w, w);
    if (big) r = 1.5707963267948966f - (r + r);
    return (f2u(x) >> 31) ? -r : r;
}

}  // namespace rtw
