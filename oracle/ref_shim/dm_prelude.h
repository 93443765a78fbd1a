// ORACLE — TEST INFRASTRUCTURE ONLY (oracle/ref_shim).
// dm_prelude.h — force-included (-include) in front of the reference's DEVICE files for the
// `libref_dm.so` build only: the six transcendental functions those files take from CUDA's libm
// are mapped to the arithmetic contract of oracle/detmath.h.  With this mapping
// the reference's own programs and the oracle must produce the SAME BITS; without it
// (`libref.so`, glibc libm) they may differ by the last-ulp differences of the two libms.
// Host-side files (main.cpp, Director.cpp: rotation matrices, camera tangent) are never built
// with this prelude.
#pragma once
#include <cmath>
#include <cstdlib>
#include <cstdio>
#include "../detmath.h"
static inline float ref_dm_powf(float x, float y) {
    if (y != 5.f) { std::fprintf(stderr, "ref_shim: powf with exponent %g (the contract defines x^5 only)\n", (double)y); std::abort(); }
    return orc::dm::pow5(x);
}
#define sinf(x) (::orc::dm::sin(x))
#define cosf(x) (::orc::dm::cos(x))
#define logf(x) (::orc::dm::log(x))
#define atan2f(y, x) (::orc::dm::atan2((y), (x)))
#define asinf(x) (::orc::dm::asin(x))
#define powf(x, y) (ref_dm_powf((x), (y)))
