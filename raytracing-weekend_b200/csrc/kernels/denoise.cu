// denoise.cu — the step AFTER the path: a stand-in for the reference's denoise pass.
//
// The reference hands its frame to the OptiX AI denoiser (Director.cpp:887-949 setup, :988-1000 invoke; LDR model
// on the gamma-encoded float4 frame).  That network lives inside the NVIDIA driver and cannot be matched; this is a
// DOCUMENTED, DIFFERENT filter — an edge-avoiding a-trous wavelet filter (Dammertz et al. 2010) on the same
// gamma-encoded frame — and it is excluded from parity with the reference.  It is NOT applied by rtw_render: the
// product's frame is the pre-denoise frame_buffer (raygen/raygen.cu:151-158); callers opt in (rtw_denoise,
// rtw_cli -denoise N).  Its own parity is against oracle/oracle.cpp orc_denoise, bit for bit: every operation is a
// correctly rounded fp32 + - * / in a fixed order (no transcendental: the colour weight is 1 / (1 + d^2 / sigma^2)).
//
//   pass i (i = 0 .. iterations-1), tap spacing s = 2^i, 5x5 taps with the B3-spline weights h = (1,4,6,4,1)/16:
//     out(p) = sum_q h(q) w(p,q) in(q) / sum_q h(q) w(p,q),   w = 1 / (1 + |in(q) - in(p)|^2 * inv_sigma2)
//   taps outside the image are skipped; alpha is copied from the centre pixel.
#include "internal.h"

namespace rtw {

namespace {

__global__ void __launch_bounds__(256) k_atrous(int Nx, int Ny, int step, float inv_sigma2, const float4* __restrict__ in,
                                                 float4* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= Nx * Ny) return;
    const int x = i % Nx, y = i / Nx;
    const float h[5] = {0.0625f, 0.25f, 0.375f, 0.25f, 0.0625f};
    const float4 c = in[i];
    float sr = 0.f, sg = 0.f, sb = 0.f, sw = 0.f;
    for (int dy = -2; dy <= 2; dy++) {
        const int qy = y + dy * step;
        if (qy < 0 || qy >= Ny) continue;
        for (int dx = -2; dx <= 2; dx++) {
            const int qx = x + dx * step;
            if (qx < 0 || qx >= Nx) continue;
            const float4 q = __ldg(in + (size_t)qy * Nx + qx);
            const float dr = q.x - c.x, dg = q.y - c.y, db = q.z - c.z;
            const float d2 = dr * dr + dg * dg + db * db;
            const float w = (h[dx + 2] * h[dy + 2]) / (1.0f + d2 * inv_sigma2);
            sr += w * q.x; sg += w * q.y; sb += w * q.z; sw += w;
        }
    }
    out[i] = make_float4(sr / sw, sg / sw, sb / sw, c.w);
}

}  // namespace

// `a` holds the input and receives the output when `iterations` is even, else `b` does: returns which.
cudaError_t launch_denoise(cudaStream_t s, int Nx, int Ny, int iterations, float sigma, float4* a, float4* b, float4** result,
                           uint64_t* launches) {
    const int n = Nx * Ny;
    const float inv_sigma2 = 1.0f / (sigma * sigma);
    float4* src = a;
    float4* dst = b;
    for (int i = 0; i < iterations; i++) {
        k_atrous<<<(n + 255) / 256, 256, 0, s>>>(Nx, Ny, 1 << i, inv_sigma2, src, dst);
        float4* t = src; src = dst; dst = t;
    }
    if (launches) *launches += (uint64_t)iterations;
    *result = src;
    return cudaGetLastError();
}

}  // namespace rtw

// ---- rtw_denoise_static: the same filter for callers that cannot call cudaMalloc (include/rtw.h) ----------------
// No counterpart in the reference.  The ping-pong buffer that
// rtw_denoise allocates per call lives here in a file-scope __device__ array: RTW_DENOISE_STATIC_MAX_PIXELS float4
// = 8 MiB, held by every GPU that loads the library until the process exits.  In HOST code `g_denoise_scratch` is
// only a stand-in for the symbol, not a device address: the address is looked up once per device, with that
// device current, by cudaGetSymbolAddress, and that pointer is what the kernel gets.  Nothing in the array has to
// start at zero: every pass writes all Nx*Ny pixels of its destination before the next pass reads them, so what
// the previous call left behind is never read.  Every call on a device shares the array: calls that overlap on the
// GPU (two streams, two host threads) race.
#include <mutex>
#include "../api/rtw_ctx.h"

static __device__ __align__(256) float4 g_denoise_scratch[RTW_DENOISE_STATIC_MAX_PIXELS];

namespace {

constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mutex;
float4* g_static_addr[kStaticMaxDevices] = {};   // device address of g_denoise_scratch on device d, once looked up

// The caller has made `device` current.
cudaError_t denoise_scratch_address(int device, float4** out) {
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_addr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_denoise_scratch);
        if (e != cudaSuccess) return e;
        g_static_addr[device] = static_cast<float4*>(p);
    }
    *out = g_static_addr[device];
    return cudaSuccess;
}

}  // namespace

extern "C" int rtw_denoise_static(rtw_ctx* c, int Nx, int Ny, int iterations, float sigma, const float* rgba_in_dev,
                                  float* rgba_out_dev, void* cuda_stream) {
    using namespace rtw;
    if (!c) return RTW_ERR_INVALID;
    if (!rgba_in_dev || !rgba_out_dev || Nx <= 0 || Ny <= 0 || iterations < 1 || iterations > 8 || !(sigma > 0.f))
        return fail(c, RTW_ERR_INVALID, "bad denoise arguments");
    const unsigned long long npix = (unsigned long long)Nx * (unsigned long long)Ny;
    if (npix > (unsigned long long)RTW_DENOISE_STATIC_MAX_PIXELS)
        return fail(c, RTW_ERR_INVALID, "rtw_denoise_static: " + std::to_string(npix) + " pixels exceed the static scratch of " +
                                            std::to_string(RTW_DENOISE_STATIC_MAX_PIXELS) + " (RTW_DENOISE_STATIC_MAX_PIXELS); use rtw_denoise");
    const char* in_b = reinterpret_cast<const char*>(rgba_in_dev);
    const char* out_b = reinterpret_cast<const char*>(rgba_out_dev);
    const size_t bytes = sizeof(float4) * (size_t)npix;
    if (in_b < out_b + bytes && out_b < in_b + bytes)
        return fail(c, RTW_ERR_INVALID, "rtw_denoise_static: input and output overlap (the input is read-only here; use rtw_denoise to filter in place)");
    if (c->device < 0 || c->device >= kStaticMaxDevices) return fail(c, RTW_ERR_INVALID, "rtw_denoise_static: device index out of range");
    RTW_CUDA(c, cudaSetDevice(c->device));
    float4* scratch = nullptr;
    RTW_CUDA(c, denoise_scratch_address(c->device, &scratch));
    cudaStream_t s = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : c->stream;
    // pass i reads what pass i-1 wrote (pass 0: the input) and the LAST pass writes the output, so the
    // destinations alternate backwards from it: output, scratch, output, ...
    const float inv_sigma2 = 1.0f / (sigma * sigma);
    const float4* src = reinterpret_cast<const float4*>(rgba_in_dev);
    float4* const out = reinterpret_cast<float4*>(rgba_out_dev);
    const int n = (int)npix;
    for (int i = 0; i < iterations; i++) {
        float4* dst = ((iterations - 1 - i) & 1) ? scratch : out;
        rtw::k_atrous<<<(n + 255) / 256, 256, 0, s>>>(Nx, Ny, 1 << i, inv_sigma2, src, dst);
        src = dst;
    }
    RTW_CUDA(c, cudaGetLastError());
    c->stats.kernel_launches += (uint64_t)iterations;
    return RTW_OK;
}
