// L2 read-bandwidth microbenchmark (denominator for the BVH-fetch roofline).
// A working set that fits the B200's L2 (default 48 MB) is read over and over with 16-byte loads
// by a persistent grid; the first pass warms L2, the timed passes are L2 hits.  Prints one JSON line.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l2_peak tools/l2_peak.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) k_read(const float4* __restrict__ p, size_t n4, int passes, float4* sink) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int k = 0; k < passes; k++) {
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
            const float4 v = __ldcg(p + i);   // L2 only: do not let L1 answer
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
    }
    if (acc.x == 123.456f) *sink = acc;   // never true: keeps the loads alive
}

int main(int argc, char** argv) {
    const double mb = argc > 1 ? atof(argv[1]) : 48.0;
    const size_t bytes = (size_t)(mb * 1024 * 1024), n4 = bytes / 16;
    float4 *d = nullptr, *sink = nullptr;
    cudaMalloc(&d, bytes); cudaMalloc(&sink, 16); cudaMemset(d, 0, bytes);
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    const int grid = prop.multiProcessorCount * 8, passes = 20;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_read<<<grid, 256>>>(d, n4, 2, sink);   // warm
    double best = 0;
    for (int rep = 0; rep < 10; rep++) {
        cudaEventRecord(e0);
        k_read<<<grid, 256>>>(d, n4, passes, sink);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double gbs = (double)bytes * passes / (ms * 1e-3) / 1e9;
        if (gbs > best) best = gbs;
    }
    printf("{\"l2_read_gbs\": %.1f, \"working_set_mb\": %.1f, \"l2_size_mb\": %.1f, \"sms\": %d, \"how\": \"float4 __ldcg reads of an L2-resident buffer, %d passes, best of 10, CUDA events\"}\n",
           best, mb, prop.l2CacheSize / 1048576.0, prop.multiProcessorCount, passes);
    return cudaGetLastError() == cudaSuccess ? 0 : 1;
}
