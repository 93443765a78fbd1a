// bake.cu — GPU-side scene flattening and texture baking: the three host-serial steps that sit
// between a scene description and the launch in the reference, as device kernels whose results are BIT-EQUAL to the
// host scene builder's (csrc/host/scene_builder.cpp), tests/test_bake.py:
//   k_invert_transforms   instance transform inversion (OptiX does it inside optixAccelBuild for the reference's
//                         instances, geometry/ioGeometryGroup.h:160-225; scene_builder.cpp make_transform): one thread
//                         per transform, float64 cofactors in the same operation order, rounded once
//   k_mt19937 + k_perlin  Perlin tables (texture/ioTexture.h:21-26 static mt19937(0) + uniform_real_distribution<float>,
//                         :137-144 permute, :181-216 getTexRec): the draw stream is produced by one thread (the twist
//                         recurrence is serial), the 256 unit vectors of a table by 256 threads, its three
//                         Fisher-Yates shuffles by three threads (a shuffle is a serial chain of swaps)
//   k_flip_rgba           image upload: vertical flip + RGB(A) -> RGBA8 (texture/ioTexture.h:247-266), one thread per texel
#include "internal.h"

namespace rtw {

namespace {

__global__ void k_invert_transforms(const float* __restrict__ o2w, int n, rtw_transform* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* m = o2w + 12 * (size_t)i;
    rtw_transform t;
    for (int k = 0; k < 12; k++) t.o2w[k] = m[k];
    double a[3][3], tr[3];
    for (int r = 0; r < 3; r++) {
        for (int c = 0; c < 3; c++) a[r][c] = (double)m[r * 4 + c];
        tr[r] = (double)m[r * 4 + 3];
    }
    const double c00 = a[1][1] * a[2][2] - a[1][2] * a[2][1];
    const double c01 = a[1][2] * a[2][0] - a[1][0] * a[2][2];
    const double c02 = a[1][0] * a[2][1] - a[1][1] * a[2][0];
    const double det = a[0][0] * c00 + a[0][1] * c01 + a[0][2] * c02;
    double inv[3][3];
    inv[0][0] = c00 / det;
    inv[1][0] = c01 / det;
    inv[2][0] = c02 / det;
    inv[0][1] = (a[0][2] * a[2][1] - a[0][1] * a[2][2]) / det;
    inv[1][1] = (a[0][0] * a[2][2] - a[0][2] * a[2][0]) / det;
    inv[2][1] = (a[0][1] * a[2][0] - a[0][0] * a[2][1]) / det;
    inv[0][2] = (a[0][1] * a[1][2] - a[0][2] * a[1][1]) / det;
    inv[1][2] = (a[0][2] * a[1][0] - a[0][0] * a[1][2]) / det;
    inv[2][2] = (a[0][0] * a[1][1] - a[0][1] * a[1][0]) / det;
    for (int r = 0; r < 3; r++) {
        for (int c = 0; c < 3; c++) t.w2o[r * 4 + c] = (float)inv[r][c];
        t.w2o[r * 4 + 3] = (float)(-(inv[r][0] * tr[0] + inv[r][1] * tr[1] + inv[r][2] * tr[2]));
    }
    out[i] = t;
}

// std::mt19937(seed) followed by libstdc++'s uniform_real_distribution<float>(0, 1): draw = gen() / 2^32 in fp32,
// a result of 1.0f replaced by the largest float below 1 (generate_canonical).  Skips `first` draws, writes `count`.
__global__ void k_mt19937(uint32_t seed, unsigned long long first, int count, float* __restrict__ out) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    __shared__ uint32_t mt[624];
    mt[0] = seed;
    for (int i = 1; i < 624; i++) mt[i] = 1812433253u * (mt[i - 1] ^ (mt[i - 1] >> 30)) + (uint32_t)i;
    int idx = 624;
    const unsigned long long total = first + (unsigned long long)count;
    for (unsigned long long k = 0; k < total; k++) {
        if (idx >= 624) {
            for (int i = 0; i < 624; i++) {
                const uint32_t y = (mt[i] & 0x80000000u) | (mt[(i + 1) % 624] & 0x7fffffffu);
                mt[i] = mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
            }
            idx = 0;
        }
        uint32_t y = mt[idx++];
        y ^= y >> 11;
        y ^= (y << 7) & 0x9d2c5680u;
        y ^= (y << 15) & 0xefc60000u;
        y ^= y >> 18;
        if (k >= first) {
            float r = (float)y / 4294967296.0f;
            if (r >= 1.0f) r = 0.99999994f;
            out[k - first] = r;
        }
    }
}

// One CTA of 256 threads per table; a table consumes 256*3 + 3*255 = 1533 draws of the stream.
__global__ void __launch_bounds__(256) k_perlin(const float* __restrict__ draws, rtw_perlin* __restrict__ out) {
    const float* d = draws + (size_t)blockIdx.x * 1533;
    rtw_perlin& p = out[blockIdx.x];
    const int i = threadIdx.x;
    {   // unit_float3(-1 + 2 r, -1 + 2 r, -1 + 2 r): the three draws of vector i in the order z, y, x (scene_builder.cpp tex_noise)
        const float z = -1 + 2 * d[3 * i + 0];
        const float y = -1 + 2 * d[3 * i + 1];
        const float x = -1 + 2 * d[3 * i + 2];
        const float l = sqrtf(x * x + y * y + z * z);
        p.ranvec[i][0] = x / l;
        p.ranvec[i][1] = y / l;
        p.ranvec[i][2] = z / l;
    }
    if (i < 3) {   // permute(), texture/ioTexture.h:137-144
        int32_t* pm = i == 0 ? p.perm_x : (i == 1 ? p.perm_y : p.perm_z);
        const float* r = d + 768 + 255 * i;
        for (int k = 0; k < 256; k++) pm[k] = k;
        int used = 0;
        for (int k = 256 - 1; k > 0; k--) {
            const int target = int(r[used++] * (k + 1));
            const int tmp = pm[k];
            pm[k] = pm[target];
            pm[target] = tmp;
        }
    }
}

__global__ void __launch_bounds__(256) k_flip_rgba(const uint8_t* __restrict__ src, int w, int h, int channels, uchar4* __restrict__ dst) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= w * h) return;
    const int x = i % w, j = i / w;
    const size_t s = ((size_t)(h - j - 1) * w + x) * channels;
    dst[i] = make_uchar4(src[s], src[s + 1], src[s + 2], channels == 4 ? src[s + 3] : (unsigned char)255);
}

}  // namespace

cudaError_t launch_invert_transforms(cudaStream_t s, const float* o2w, int n, rtw_transform* out) {
    k_invert_transforms<<<(n + 127) / 128, 128, 0, s>>>(o2w, n, out);
    return cudaGetLastError();
}
cudaError_t launch_perlin_tables(cudaStream_t s, uint32_t seed, int first_table, int n_tables, float* draws, rtw_perlin* out) {
    k_mt19937<<<1, 32, 0, s>>>(seed, (unsigned long long)first_table * 1533ull, n_tables * 1533, draws);
    k_perlin<<<n_tables, 256, 0, s>>>(draws, out);
    return cudaGetLastError();
}
cudaError_t launch_flip_rgba(cudaStream_t s, const uint8_t* src, int w, int h, int channels, uchar4* dst) {
    k_flip_rgba<<<(w * h + 255) / 256, 256, 0, s>>>(src, w, h, channels, dst);
    return cudaGetLastError();
}

}  // namespace rtw
