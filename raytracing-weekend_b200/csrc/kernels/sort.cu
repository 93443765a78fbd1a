// sort.cu — stable LSD radix sort of (key, 32-bit value) pairs: the 63-bit Morton keys of the LBVH build (step 3 of
// SURVEY.md §7 step 4).  8 bits per pass; per pass
//
//   k_sort_tile_hist   digit histogram of every 4096-key tile -> counts[digit][tile], and the digit totals of the pass
//   k_sort_scan        one CTA per DIGIT: exclusive scan of its row of tile counts (coalesced) + the digit's global
//                      base (sum of the totals of the lower digits) -> the output position of (digit, tile)
//   k_sort_scatter     a warp owns a contiguous run of 16 x 32 keys: ranks equal digits with __match_any_sync in
//                      index order, warps are ordered by a per-digit scan over the per-warp histograms, then scatters
//
// Stability: tiles by the scan, warps by the scan over warps, keys inside a warp by index — the one stable order
// (bit-equal to std::stable_sort in the oracle's mirror).  History (profiles/r02/build_r02f.md): round 1 scanned all
// 256 x tiles counters on ONE CTA (231 us per pass at 2^20 keys, 58 % of the whole build); a one-sweep variant with
// decoupled look-back was tried next and is slower at these sizes than three plain launches (40 us per pass at 2^20:
// every tile of a 2^20..2^21-key array is resident at once, so no tile finds a finished predecessor and each walks
// the whole chain of aggregates).
//
// Bound: HBM.  Algorithmic bytes per pass = n * sizeof(key) (histogram) + n * (sizeof(key) + 4) * 2 (scatter).
#include "internal.h"

namespace rtw {

namespace {

constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kItems = 16;                       // keys per thread
constexpr int kTile = kSortThreads * kItems;     // 4096 keys per tile
constexpr int kMaxPasses = 8;

// scratch layout (uint32 words): [0, 2048) digit totals [pass][256]; [2048, 2048 + 256 * tiles) counts [digit][tile]
constexpr int kOffCounts = kMaxPasses * 256;

template <class K>
__device__ __forceinline__ uint32_t digit_of(K key, int shift) { return (uint32_t)(key >> shift) & 0xffu; }

template <class K>
__global__ void __launch_bounds__(kSortThreads) k_sort_tile_hist(const K* __restrict__ keys, int n, int shift,
                                                                  uint32_t* __restrict__ counts, int tiles,
                                                                  uint32_t* __restrict__ totals) {
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
    const uint32_t c = hist[threadIdx.x];
    counts[(size_t)threadIdx.x * tiles + blockIdx.x] = c;
    if (c) atomicAdd(totals + threadIdx.x, c);
}

// CTA d: counts[d][0..tiles) -> exclusive prefix + base of digit d.  Thread t scans `per` consecutive tiles.
__global__ void __launch_bounds__(kSortThreads) k_sort_scan(uint32_t* __restrict__ counts, int tiles, const uint32_t* __restrict__ totals) {
    __shared__ uint32_t s_warp[kSortWarps];
    __shared__ uint32_t s_base;
    const int d = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // base of the digit: the keys of all lower digits
    uint32_t v = (int)threadIdx.x < d ? totals[threadIdx.x] : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s_warp[warp] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t b = 0;
        for (int w = 0; w < kSortWarps; w++) b += s_warp[w];
        s_base = b;
    }
    __syncthreads();
    uint32_t* row = counts + (size_t)d * tiles;
    const int per = (tiles + kSortThreads - 1) / kSortThreads;
    // (a thread's `per` tiles are consecutive; for tiles <= 256 this is one coalesced load per thread)
    const int begin = threadIdx.x * per, end = min(begin + per, tiles);
    uint32_t sum = 0;
    for (int i = begin; i < end; i++) sum += row[i];
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += u;
    }
    __syncthreads();   // s_warp is reused
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t run = s_base + inc - sum;
    for (int w = 0; w < warp; w++) run += s_warp[w];
    for (int i = begin; i < end; i++) {
        const uint32_t c = row[i];
        row[i] = run;
        run += c;
    }
}

template <class K>
__global__ void __launch_bounds__(kSortThreads) k_sort_scatter(const K* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                                K* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int n,
                                                                int shift, const uint32_t* __restrict__ scanned, int tiles) {
    __shared__ uint32_t warp_hist[kSortWarps][256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < kSortWarps * 256; i += kSortThreads) (&warp_hist[0][0])[i] = 0;
    __syncthreads();
    // a warp owns a contiguous run of kItems * 32 keys, in index order
    const int base = blockIdx.x * kTile + warp * (kItems * 32);
    K key[kItems];
    uint32_t rank[kItems];
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < kItems; r++) {
        const int i = base + r * 32 + lane;
        key[r] = i < n ? keys_in[i] : (K)0;
    }
#pragma unroll
    for (int r = 0; r < kItems; r++) {
        const int i = base + r * 32 + lane;
        const bool valid = i < n;
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
    {   // thread d: exclusive scan of digit d over the warps + the position of (digit, tile)
        const int d = threadIdx.x;
        uint32_t run = scanned[(size_t)d * tiles + blockIdx.x];
#pragma unroll
        for (int w = 0; w < kSortWarps; w++) {
            const uint32_t c = warp_hist[w][d];
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
            vals_out[pos] = vals_in[i];
        }
    }
}

template <class K>
cudaError_t sort_pairs_t(cudaStream_t s, K* keys[2], uint32_t* vals[2], int n, uint32_t* scratch, uint64_t* launches, int passes) {
    if (n <= 0) return cudaSuccess;
    const int max_passes = (int)sizeof(K);
    if (passes < 1 || passes > max_passes) passes = max_passes;
    const int tiles = (n + kTile - 1) / kTile;
    cudaError_t e = cudaMemsetAsync(scratch, 0, sizeof(uint32_t) * kOffCounts, s);   // the digit totals of every pass
    if (e != cudaSuccess) return e;
    uint32_t* counts = scratch + kOffCounts;
    int cur = 0;
    for (int pass = 0; pass < passes; pass++) {
        const int shift = 8 * pass;
        uint32_t* totals = scratch + 256 * pass;
        k_sort_tile_hist<K><<<tiles, kSortThreads, 0, s>>>(keys[cur], n, shift, counts, tiles, totals);
        k_sort_scan<<<256, kSortThreads, 0, s>>>(counts, tiles, totals);
        k_sort_scatter<K><<<tiles, kSortThreads, 0, s>>>(keys[cur], vals[cur], keys[cur ^ 1], vals[cur ^ 1], n, shift, counts, tiles);
        cur ^= 1;
    }
    if (launches) *launches += 3 * (uint64_t)passes;
    // an even number of passes (the build's 8): the result is back in buffer 0; odd: in buffer 1
    return cudaGetLastError();
}

}  // namespace

size_t sort_scratch_words(int n) {
    const int tiles = (n + kTile - 1) / kTile;
    return (size_t)kOffCounts + (size_t)256 * (size_t)(tiles > 0 ? tiles : 1);
}

cudaError_t sort_pairs(cudaStream_t s, unsigned long long* keys[2], uint32_t* vals[2], int n, uint32_t* scratch,
                       uint64_t* launches, int passes) {
    return sort_pairs_t<unsigned long long>(s, keys, vals, n, scratch, launches, passes);
}

}  // namespace rtw
