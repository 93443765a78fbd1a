// ORACLE — TEST INFRASTRUCTURE ONLY.  Never linked into, imported by or called
// from the product library (raytracing-weekend_b200/).  See oracle/README.md.
//
// vecmath.h — CPU restatement of the float3 helpers the reference takes from
// the OptiX SDK 8.0.0 `sutil/vec_math.h` (pinned by RestOfLife.vcxproj:112; the
// header itself is NOT under /root/reference → "parity unpinned").
// Restated from the published SDK definitions:
//   normalize(v) = v * (1/sqrtf(dot(v,v)));  length = sqrtf(dot);  v / s = v * (1/s);
//   reflect(i,n) = i - 2*n*dot(n,i);  fmaxf(float3) = max of components.
// Build with -ffp-contract=off so a*b+c is two roundings, exactly as written.
#pragma once
#include <cmath>
#include <cstdint>

namespace orc {

struct V3 { float x, y, z; };

static inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
static inline V3 v3(float s) { return V3{s, s, s}; }
static inline V3 v3(const float* p) { return V3{p[0], p[1], p[2]}; }

static inline V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
static inline V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
static inline V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
static inline V3 operator/(V3 a, V3 b) { return v3(a.x / b.x, a.y / b.y, a.z / b.z); }
// sutil: float3 / float multiplies by the reciprocal
static inline V3 operator/(V3 a, float s) { float inv = 1.0f / s; return a * inv; }
static inline V3 operator+(V3 a, float s) { return v3(a.x + s, a.y + s, a.z + s); }
static inline V3 operator-(V3 a, float s) { return v3(a.x - s, a.y - s, a.z - s); }
static inline V3& operator+=(V3& a, V3 b) { a = a + b; return a; }
static inline V3& operator*=(V3& a, V3 b) { a = a * b; return a; }
static inline V3& operator*=(V3& a, float s) { a = a * s; return a; }
static inline V3& operator/=(V3& a, float s) { float inv = 1.0f / s; a = a * inv; return a; }

static inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline V3 cross(V3 a, V3 b) {
    return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline float length(V3 a) { return sqrtf(dot(a, a)); }
static inline V3 normalize(V3 a) { float inv = 1.0f / sqrtf(dot(a, a)); return a * inv; }
static inline V3 reflect(V3 i, V3 n) { return i - 2.0f * n * dot(n, i); }
static inline float fmaxf3(V3 a) { return fmaxf(fmaxf(a.x, a.y), a.z); }
static inline float fminf3(V3 a) { return fminf(fminf(a.x, a.y), a.z); }
static inline V3 vmin(V3 a, V3 b) { return v3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
static inline V3 vmax(V3 a, V3 b) { return v3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }

// CUDA math_constants.h values used by the reference
static const float kPi    = 3.141592654f;   // CUDART_PI_F
static const float kPiO2  = 1.570796327f;   // CUDART_PIO2_F
static const float kInvPi = 0.318309886183790671538f;  // sutil M_1_PIf

// Row-major 3x4 affine helpers; evaluation order ((m0*x + m1*y) + m2*z) + m3.
static inline V3 xform_point(const float* m, V3 p) {
    return v3(m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3],
              m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7],
              m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]);
}
static inline V3 xform_vector(const float* m, V3 d) {
    return v3(m[0] * d.x + m[1] * d.y + m[2] * d.z,
              m[4] * d.x + m[5] * d.y + m[6] * d.z,
              m[8] * d.x + m[9] * d.y + m[10] * d.z);
}
// normal object→world = (world→object linear part)^T * n
static inline V3 xform_normal(const float* w2o, V3 n) {
    return v3(w2o[0] * n.x + w2o[4] * n.y + w2o[8] * n.z,
              w2o[1] * n.x + w2o[5] * n.y + w2o[9] * n.z,
              w2o[2] * n.x + w2o[6] * n.y + w2o[10] * n.z);
}

}  // namespace orc
