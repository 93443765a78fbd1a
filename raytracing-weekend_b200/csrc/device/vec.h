// vec.h — fp32 vector helpers for the device code.
//
// Functions are __host__ __device__ so the same per-thread bodies can be driven
// by tests/emu (a test-only host harness); the product only ever runs them in
// CUDA kernels.  Arithmetic follows the OptiX SDK sutil/vec_math.h semantics the
// reference was written against (v / s multiplies by 1/s; normalize multiplies by
// 1/sqrtf(dot)), and every file is compiled with -fmad=false so a*b+c rounds
// twice unless fmaf() is written explicitly.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define RTW_HD __host__ __device__ __forceinline__
#define RTW_HDN __host__ __device__
#else
#define RTW_HD inline
#define RTW_HDN
#endif

namespace rtw {

RTW_HD float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
RTW_HD float3 f3(float s) { return make_float3(s, s, s); }
RTW_HD float3 f3(float4 v) { return make_float3(v.x, v.y, v.z); }

RTW_HD float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
RTW_HD float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
RTW_HD float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
RTW_HD float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
RTW_HD float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
RTW_HD float3 operator*(float s, float3 a) { return f3(s * a.x, s * a.y, s * a.z); }
RTW_HD float3 operator/(float3 a, float3 b) { return f3(a.x / b.x, a.y / b.y, a.z / b.z); }
RTW_HD float3 operator/(float3 a, float s) { float inv = 1.0f / s; return a * inv; }
RTW_HD float3 operator+(float3 a, float s) { return f3(a.x + s, a.y + s, a.z + s); }
RTW_HD float3 operator-(float3 a, float s) { return f3(a.x - s, a.y - s, a.z - s); }

RTW_HD float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
RTW_HD float3 cross(float3 a, float3 b) {
    return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
RTW_HD float length(float3 a) { return sqrtf(dot(a, a)); }
RTW_HD float3 normalize(float3 a) { float inv = 1.0f / sqrtf(dot(a, a)); return a * inv; }
RTW_HD float3 reflect(float3 i, float3 n) { return i - 2.0f * n * dot(n, i); }
RTW_HD float max3(float3 a) { return fmaxf(fmaxf(a.x, a.y), a.z); }
RTW_HD float min3(float3 a) { return fminf(fminf(a.x, a.y), a.z); }
RTW_HD float3 vmin(float3 a, float3 b) { return f3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
RTW_HD float3 vmax(float3 a, float3 b) { return f3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }

// bit casts
RTW_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
RTW_HD float u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}

#define RTW_PI      3.141592654f              /* CUDART_PI_F   */
#define RTW_PIO2    1.570796327f              /* CUDART_PIO2_F */
#define RTW_INV_PI  0.318309886183790671538f  /* sutil M_1_PIf */
#define RTW_FLT_MAX 3.402823466e+38f

// Row-major 3x4 affine stored as three float4 rows; ((m0*x + m1*y) + m2*z) + m3.
RTW_HD float3 xform_point(float4 r0, float4 r1, float4 r2, float3 p) {
    return f3(r0.x * p.x + r0.y * p.y + r0.z * p.z + r0.w,
              r1.x * p.x + r1.y * p.y + r1.z * p.z + r1.w,
              r2.x * p.x + r2.y * p.y + r2.z * p.z + r2.w);
}
RTW_HD float3 xform_vector(float4 r0, float4 r1, float4 r2, float3 d) {
    return f3(r0.x * d.x + r0.y * d.y + r0.z * d.z,
              r1.x * d.x + r1.y * d.y + r1.z * d.z,
              r2.x * d.x + r2.y * d.y + r2.z * d.z);
}
// normal object→world = (world→object linear part)^T n
RTW_HD float3 xform_normal(float4 r0, float4 r1, float4 r2, float3 n) {
    return f3(r0.x * n.x + r1.x * n.y + r2.x * n.z,
              r0.y * n.x + r1.y * n.y + r2.y * n.z,
              r0.z * n.x + r1.z * n.y + r2.z * n.z);
}

}  // namespace rtw
