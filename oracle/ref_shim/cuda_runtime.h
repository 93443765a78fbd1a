// ORACLE — TEST INFRASTRUCTURE ONLY (oracle/ref_shim).
// cuda_runtime.h — the CUDA runtime calls the reference's host code makes, with "device
// memory" being host memory (cudaMalloc = malloc, cudaMemcpy = memcpy) and texture objects
// being plain RGBA8 arrays sampled by refshim::tex2d (ref_runtime.cpp).
#pragma once
#include "vector_types.h"
#include "texture_types.h"

typedef enum cudaError { cudaSuccess = 0, cudaErrorInvalidValue = 1 } cudaError_t;
typedef struct CUstream_st* cudaStream_t;
typedef cudaStream_t CUstream;
enum cudaMemcpyKind { cudaMemcpyHostToHost = 0, cudaMemcpyHostToDevice = 1, cudaMemcpyDeviceToHost = 2, cudaMemcpyDeviceToDevice = 3 };

cudaError_t cudaMalloc(void** p, size_t n);
cudaError_t cudaFree(void* p);
cudaError_t cudaMemcpy(void* dst, const void* src, size_t n, cudaMemcpyKind kind);
cudaError_t cudaStreamCreate(cudaStream_t* s);
cudaError_t cudaStreamDestroy(cudaStream_t s);
cudaError_t cudaSetDevice(int);

enum cudaChannelFormatKind { cudaChannelFormatKindSigned = 0, cudaChannelFormatKindUnsigned = 1, cudaChannelFormatKindFloat = 2 };
struct cudaChannelFormatDesc { int x, y, z, w; cudaChannelFormatKind f; };
static inline cudaChannelFormatDesc cudaCreateChannelDesc(int x, int y, int z, int w, cudaChannelFormatKind f) { return cudaChannelFormatDesc{x, y, z, w, f}; }
struct cudaArray { int width, height; unsigned char* rgba; };
typedef cudaArray* cudaArray_t;
cudaError_t cudaMallocArray(cudaArray_t* a, const cudaChannelFormatDesc* d, size_t w, size_t h);
cudaError_t cudaMemcpy2DToArray(cudaArray_t dst, size_t wOffset, size_t hOffset, const void* src, size_t spitch, size_t widthBytes, size_t height,
                                cudaMemcpyKind kind);
enum cudaResourceType { cudaResourceTypeArray = 0 };
struct cudaResourceDesc { cudaResourceType resType; union { struct { cudaArray_t array; } array; } res; };
enum cudaTextureAddressMode { cudaAddressModeWrap = 0, cudaAddressModeClamp = 1 };
enum cudaTextureFilterMode { cudaFilterModePoint = 0, cudaFilterModeLinear = 1 };
enum cudaTextureReadMode { cudaReadModeElementType = 0, cudaReadModeNormalizedFloat = 1 };
struct cudaTextureDesc {
    cudaTextureAddressMode addressMode[3]; cudaTextureFilterMode filterMode; cudaTextureReadMode readMode; int sRGB; float borderColor[4];
    int normalizedCoords; unsigned int maxAnisotropy; cudaTextureFilterMode mipmapFilterMode; float mipmapLevelBias, minMipmapLevelClamp,
    maxMipmapLevelClamp;
};
struct cudaResourceViewDesc { int unused; };
cudaError_t cudaCreateTextureObject(cudaTextureObject_t* out, const cudaResourceDesc* res, const cudaTextureDesc* tex, const cudaResourceViewDesc* view);
