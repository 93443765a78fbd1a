// ORACLE — TEST INFRASTRUCTURE ONLY (oracle/ref_shim).  sutil::CUDAOutputBuffer as a host
// vector: "device" memory is host memory in the stand-in.
#pragma once
#include <vector>
#include "Exception.h"
namespace sutil {
enum class CUDAOutputBufferType { CUDA_DEVICE = 0, GL_INTEROP = 1, ZERO_COPY = 2, CUDA_P2P = 3 };
template <typename PIXEL_FORMAT>
class CUDAOutputBuffer {
public:
    CUDAOutputBuffer(CUDAOutputBufferType, int width, int height) : m_pixels((size_t)width * height) {}
    void setStream(CUstream) {}
    PIXEL_FORMAT* map() { return m_pixels.data(); }
    void unmap() {}
    PIXEL_FORMAT* getHostPointer() { return m_pixels.data(); }
private:
    std::vector<PIXEL_FORMAT> m_pixels;
};
}
