// ORACLE — TEST INFRASTRUCTURE ONLY (oracle/ref_shim).
// OptiX SDK 8.0.0 SDK/sutil/vec_math.h is not under /root/reference (SURVEY.md §8c,
// "parity unpinned" for this header).  Restated from its published definitions, the subset
// the reference uses:  normalize(v) = v * (1 / sqrtf(dot(v, v)));  v / s = v * (1 / s);
// reflect(i, n) = i - 2 * n * dot(n, i);  fmaxf(float3) = max of the components;
// clamp(f, a, b) = fmaxf(a, fminf(f, b)).  Written independently of oracle/vecmath.h.
#pragma once
#include "../vector_types.h"
#include <cmath>
#include <cstdlib>

#ifndef M_PIf
#define M_PIf 3.14159265358979323846f
#endif
#ifndef M_PI_2f
#define M_PI_2f 1.57079632679489661923f
#endif
#ifndef M_1_PIf
#define M_1_PIf 0.318309886183790671538f
#endif

// CUDA provides min/max overloads in the global namespace
static inline float min(float a, float b) { return fminf(a, b); }
static inline float max(float a, float b) { return fmaxf(a, b); }
static inline int min(int a, int b) { return a < b ? a : b; }
static inline int max(int a, int b) { return a > b ? a : b; }
static inline unsigned min(unsigned a, unsigned b) { return a < b ? a : b; }
static inline unsigned max(unsigned a, unsigned b) { return a > b ? a : b; }

static inline float clamp(float f, float a, float b) { return fmaxf(a, fminf(f, b)); }
static inline int clamp(int f, int a, int b) { return max(a, min(f, b)); }

// ---- float2 ----
static inline float2 make_float2(float s) { return make_float2(s, s); }
static inline float2 operator+(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
static inline float2 operator-(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
static inline float2 operator*(float2 a, float s) { return make_float2(a.x * s, a.y * s); }
static inline float2 operator*(float s, float2 a) { return make_float2(s * a.x, s * a.y); }

// ---- float3 ----
#ifdef REF_SHIM_LTR_ARGS
static inline float3 ref_make_float3_1(float s) { return float3{s, s, s}; }
static inline float3 ref_make_float3_1(const float4& v) { return float3{v.x, v.y, v.z}; }
#else
static inline float3 make_float3(float s) { return make_float3(s, s, s); }
static inline float3 make_float3(const float4& v) { return make_float3(v.x, v.y, v.z); }
#endif
static inline float3 operator-(const float3& a) { return make_float3(-a.x, -a.y, -a.z); }
static inline float3 operator+(const float3& a, const float3& b) { return make_float3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline float3 operator+(const float3& a, float b) { return make_float3(a.x + b, a.y + b, a.z + b); }
static inline float3 operator+(float a, const float3& b) { return make_float3(a + b.x, a + b.y, a + b.z); }
static inline void operator+=(float3& a, const float3& b) { a.x += b.x; a.y += b.y; a.z += b.z; }
static inline float3 operator-(const float3& a, const float3& b) { return make_float3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline float3 operator-(const float3& a, float b) { return make_float3(a.x - b, a.y - b, a.z - b); }
static inline float3 operator-(float a, const float3& b) { return make_float3(a - b.x, a - b.y, a - b.z); }
static inline void operator-=(float3& a, const float3& b) { a.x -= b.x; a.y -= b.y; a.z -= b.z; }
static inline float3 operator*(const float3& a, const float3& b) { return make_float3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline float3 operator*(const float3& a, float s) { return make_float3(a.x * s, a.y * s, a.z * s); }
static inline float3 operator*(float s, const float3& a) { return make_float3(s * a.x, s * a.y, s * a.z); }
static inline void operator*=(float3& a, const float3& s) { a.x *= s.x; a.y *= s.y; a.z *= s.z; }
static inline void operator*=(float3& a, float s) { a.x *= s; a.y *= s; a.z *= s; }
static inline float3 operator/(const float3& a, const float3& b) { return make_float3(a.x / b.x, a.y / b.y, a.z / b.z); }
static inline float3 operator/(const float3& a, float s) { float inv = 1.0f / s; return a * inv; }
static inline float3 operator/(float s, const float3& a) { return make_float3(s / a.x, s / a.y, s / a.z); }
static inline void operator/=(float3& a, float s) { float inv = 1.0f / s; a *= inv; }

static inline float3 fminf(const float3& a, const float3& b) { return make_float3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
static inline float fminf(const float3& a) { return fminf(fminf(a.x, a.y), a.z); }
static inline float3 fmaxf(const float3& a, const float3& b) { return make_float3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }
static inline float fmaxf(const float3& a) { return fmaxf(fmaxf(a.x, a.y), a.z); }

static inline float dot(const float3& a, const float3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline float3 cross(const float3& a, const float3& b) {
    return make_float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline float length(const float3& v) { return sqrtf(dot(v, v)); }
static inline float3 normalize(const float3& v) { float invLen = 1.0f / sqrtf(dot(v, v)); return v * invLen; }
static inline float3 reflect(const float3& i, const float3& n) { return i - 2.0f * n * dot(n, i); }
static inline float3 clamp(const float3& v, float a, float b) { return make_float3(clamp(v.x, a, b), clamp(v.y, a, b), clamp(v.z, a, b)); }
static inline float3 expf(const float3& v) { return make_float3(::expf(v.x), ::expf(v.y), ::expf(v.z)); }

// ---- float4 ----
static inline float4 make_float4(float s) { return make_float4(s, s, s, s); }
static inline float4 make_float4(const float3& a, float w) { return make_float4(a.x, a.y, a.z, w); }
static inline float4 operator+(const float4& a, const float4& b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
static inline float4 operator*(const float4& a, float s) { return make_float4(a.x * s, a.y * s, a.z * s, a.w * s); }
static inline float4 operator*(float s, const float4& a) { return make_float4(s * a.x, s * a.y, s * a.z, s * a.w); }

template <typename IntegerType>
static inline IntegerType roundUp(IntegerType x, IntegerType y) { return ((x + y - 1) / y) * y; }
