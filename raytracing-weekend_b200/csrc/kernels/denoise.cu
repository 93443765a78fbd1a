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
