// ORACLE — TEST INFRASTRUCTURE ONLY (oracle/ref_shim).
// OptiX SDK 8.0.0 SDK/sutil/Matrix.h is not under /root/reference.  Restated from its
// published definition, the subset the reference uses (geometry/ioTransform.h,
// scene/ioScene.h:546-548,718-720,942-943): row-major 4x4, construction from 16 floats,
// getData(), operator*= as a right-multiply with the k-ascending dot product.
#pragma once
#include "vec_math.h"
namespace sutil {
class Matrix4x4 {
public:
    Matrix4x4() { for (int i = 0; i < 16; i++) m_data[i] = 0.f; }
    explicit Matrix4x4(const float* d) { for (int i = 0; i < 16; i++) m_data[i] = d[i]; }
    float* getData() { return m_data; }
    const float* getData() const { return m_data; }
    float& operator[](unsigned i) { return m_data[i]; }
    float operator[](unsigned i) const { return m_data[i]; }
    static Matrix4x4 identity() {
        Matrix4x4 m;
        m.m_data[0] = m.m_data[5] = m.m_data[10] = m.m_data[15] = 1.f;
        return m;
    }
    static Matrix4x4 scale(const float3& s) {
        Matrix4x4 m = identity();
        m.m_data[0] = s.x; m.m_data[5] = s.y; m.m_data[10] = s.z;
        return m;
    }
private:
    float m_data[16];
};
inline Matrix4x4 operator*(const Matrix4x4& a, const Matrix4x4& b) {
    Matrix4x4 r;
    for (unsigned i = 0; i < 4; ++i)
        for (unsigned j = 0; j < 4; ++j) {
            float sum = 0.0f;
            for (unsigned k = 0; k < 4; ++k) sum += a[i * 4 + k] * b[k * 4 + j];
            r[i * 4 + j] = sum;
        }
    return r;
}
inline Matrix4x4& operator*=(Matrix4x4& a, const Matrix4x4& b) { a = a * b; return a; }
}  // namespace sutil
