// sort.cu — stable LSD radix sort of (63-bit Morton key, primitive index) pairs.
//
// Step 3 of the LBVH build.  Eight passes of 8 bits; each
// pass is histogram → exclusive scan → stable scatter.  Stability inside a block
// comes from processing keys in index order: a warp owns a contiguous run of
// kItems*32 keys and ranks equal digits with __match_any_sync, warps are ordered
// by a per-digit exclusive scan over the per-warp histograms.
//
// Bound: HBM.  Algorithmic bytes per pass = n * (8 + 4) * 2 (read + write of the
// pairs) + n * 8 for the histogram read; 8 passes.
#include "internal.h"

namespace rtw {

namespace {

constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kItems = 8;                        // keys per thread
constexpr int kTile = kSortThreads * kItems;     // 2048 keys per block

__device__ __forceinline__ uint32_t digit_of(unsigned long long key, int shift) { return (uint32_t)(key >> shift) & 0xffu; }

__global__ void __launch_bounds__(kSortThreads) k_sort_hist(const unsigned long long* __restrict__ keys, int n, int shift,
                                                             uint32_t* __restrict__ block_hist, int num_blocks) {
    __shared__ uint32_t hist[256];
    hist[threadIdx.x] = 0;
    __syncthreads();
    const int base = blockIdx.x * kTile;
#pragma unroll
    for (int r = 0; r < kItems; r++) {
        const int i = base + r * kSortThreads + threadIdx.x;
        if (i < n) atomicAdd(&hist[digit_of(keys[i], shift)], 1u);
    }
    __syncthreads();
    block_hist[threadIdx.x * num_blocks + blockIdx.x] = hist[threadIdx.x];
}

// Exclusive scan of `count` uint32 in place, one block of 1024 threads.
__global__ void __launch_bounds__(1024) k_sort_scan(uint32_t* __restrict__ data, int count) {
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry;
    const int per = (count + 1023) / 1024;
    const int begin = threadIdx.x * per;
    const int end = min(begin + per, count);
    uint32_t sum = 0;
    for (int i = begin; i < end; i++) sum += data[i];
    // block exclusive scan of the per-thread sums
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t ws = warp_sums[lane];
        uint32_t winc = ws;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t v = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += v;
        }
        warp_sums[lane] = winc - ws;
        if (lane == 31) carry = winc;
    }
    __syncthreads();
    uint32_t run = warp_sums[warp] + inc - sum;
    for (int i = begin; i < end; i++) {
        uint32_t v = data[i];
        data[i] = run;
        run += v;
    }
    (void)carry;
}

__global__ void __launch_bounds__(kSortThreads) k_sort_scatter(const unsigned long long* __restrict__ keys_in,
                                                                const uint32_t* __restrict__ vals_in,
                                                                unsigned long long* __restrict__ keys_out,
                                                                uint32_t* __restrict__ vals_out, int n, int shift,
                                                                const uint32_t* __restrict__ scanned_hist, int num_blocks) {
    __shared__ uint32_t warp_hist[kSortWarps][256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < kSortWarps * 256; i += kSortThreads) (&warp_hist[0][0])[i] = 0;
    __syncthreads();
    const int base = blockIdx.x * kTile + warp * (kItems * 32);
    unsigned long long key[kItems];
    uint32_t val[kItems];
    uint32_t rank[kItems];
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < kItems; r++) {
        const int i = base + r * 32 + lane;
        const bool valid = i < n;
        key[r] = valid ? keys_in[i] : 0ull;
        val[r] = valid ? vals_in[i] : 0u;
        // invalid lanes get a unique pseudo digit so they never match a valid one
        const uint32_t d = valid ? digit_of(key[r], shift) : (0x100u | (uint32_t)lane);
        const uint32_t mask = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(mask) - 1;
        uint32_t pre = 0;
        if (valid && lane == leader) {
            pre = warp_hist[warp][d];
            warp_hist[warp][d] = pre + __popc(mask);
        }
        pre = __shfl_sync(0xffffffffu, pre, leader);
        rank[r] = pre + __popc(mask & lt_mask);
        __syncwarp();
    }
    __syncthreads();
    {   // per digit: exclusive scan over the warps + global base of (digit, block)
        const int d = threadIdx.x;
        uint32_t run = scanned_hist[d * num_blocks + blockIdx.x];
#pragma unroll
        for (int w = 0; w < kSortWarps; w++) {
            uint32_t c = warp_hist[w][d];
            warp_hist[w][d] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kItems; r++) {
        const int i = base + r * 32 + lane;
        if (i < n) {
            const uint32_t pos = warp_hist[warp][digit_of(key[r], shift)] + rank[r];
            keys_out[pos] = key[r];
            vals_out[pos] = val[r];
        }
    }
}

}  // namespace

int sort_num_blocks(int n) { return (n + kTile - 1) / kTile; }

cudaError_t sort_pairs(cudaStream_t s, unsigned long long* keys[2], uint32_t* vals[2], int n, uint32_t* hist,
                       uint64_t* launches) {
    if (n <= 0) return cudaSuccess;
    const int nb = sort_num_blocks(n);
    int cur = 0;
    for (int pass = 0; pass < 8; pass++) {
        const int shift = pass * 8;
        // NOTE: k_sort_hist reads keys in a different thread→key mapping than the scatter
        // kernel; only the per-block digit counts matter and the tile is the same.
        k_sort_hist<<<nb, kSortThreads, 0, s>>>(keys[cur], n, shift, hist, nb);
        k_sort_scan<<<1, 1024, 0, s>>>(hist, 256 * nb);
        k_sort_scatter<<<nb, kSortThreads, 0, s>>>(keys[cur], vals[cur], keys[cur ^ 1], vals[cur ^ 1], n, shift, hist, nb);
        if (launches) *launches += 3;
        cur ^= 1;
    }
    // 8 passes: the result is back in buffer 0
    return cudaGetLastError();
}

}  // namespace rtw
