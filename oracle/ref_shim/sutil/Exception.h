// ORACLE — TEST INFRASTRUCTURE ONLY (oracle/ref_shim).  The SDK's error-check macros:
// every stand-in call returns success, a failure aborts with the call text.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include "../optix.h"
namespace sutil {
class Exception : public std::runtime_error {
public:
    explicit Exception(const char* msg) : std::runtime_error(msg) {}
};
}
#define CUDA_CHECK(call)  do { if ((call) != cudaSuccess) throw sutil::Exception("CUDA call failed: " #call); } while (0)
#define OPTIX_CHECK(call) do { if ((call) != OPTIX_SUCCESS) throw sutil::Exception("OptiX call failed: " #call); } while (0)
#define CUDA_SYNC_CHECK() do { } while (0)
static char LOG[2048];
static size_t LOG_SIZE = sizeof(LOG);
#define OPTIX_CHECK_LOG(call) do { LOG_SIZE = sizeof(LOG); if ((call) != OPTIX_SUCCESS) throw sutil::Exception("OptiX call failed: " #call); (void)LOG; } while (0)
