// ORACLE — TEST INFRASTRUCTURE ONLY (oracle/ref_shim).  sutil::Aabb: only named inside
// `#if 0` / comments of the reference (geometry/ioGeometryGroup.h); declared for completeness.
#pragma once
#include "vec_math.h"
namespace sutil {
class Aabb {
public:
    float3 m_min{1e37f, 1e37f, 1e37f}, m_max{-1e37f, -1e37f, -1e37f};
    void include(const float3& a, const float3& b) { m_min = fminf(m_min, a); m_max = fmaxf(m_max, b); }
};
}
