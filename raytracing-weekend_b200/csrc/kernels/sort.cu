// sort.cu — stable LSD radix sort of (63-bit Morton key, primitive index) pairs, one sweep per digit.
//
// Step 3 of the LBVH build.  Eight passes of 8 bits.  Round 1 ran histogram → single-CTA
// scan → scatter per pass (24 launches, every pass read the keys twice, and the scan of 256 x tiles counters on
// ONE CTA was the longest kernel of the build at 2^20 primitives).  Now ("onesweep", Adinets & Merrill 2022):
//
//   k_sort_hist_all   ONE pass over the keys: the 8 digit histograms of the whole array (shared-memory atomics,
//                     then 8 x 256 global atomics per CTA)
//   k_sort_onesweep   per digit: a CTA takes the next tile (atomic ticket, so tiles start in order), ranks its keys
//                     (per-warp __match_any_sync ranking + a scan over the warps, as before), publishes the tile's
//                     per-digit counts, finds the counts of all earlier tiles by DECOUPLED LOOK-BACK (a tile's
//                     status word per digit is `tag | count`: tag says aggregate-only or inclusive prefix; the tag
//                     carries the pass number so the words need no clearing between passes) and scatters
//
// = 1 memset + 9 launches, keys read 9 times and written 8 times.  Stability: tiles are ordered by the look-back,
// warps inside a tile by the scan over the per-warp histograms, keys inside a warp by index — the output is the
// one stable order, bit-equal to round 1's (and to std::stable_sort in the oracle's mirror).
//
// Bound: HBM.  Algorithmic bytes = n * 8 (histogram pass) + 8 passes * n * (8 + 4) * 2.
#include "internal.h"

namespace rtw {

namespace {

constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kItems = 16;                       // keys per thread
constexpr int kTile = kSortThreads * kItems;     // 4096 keys per tile
constexpr int kPasses = 8;

// scratch layout (uint32 words): [0, 2048) digit histograms [pass][256]; [2048, 2056) tile tickets per pass;
// [2064, 2064 + 256 * tiles) status words [tile][256]
constexpr int kOffTicket = kPasses * 256;
constexpr int kOffStatus = kOffTicket + 16;
constexpr uint32_t kCountMask = (1u << 27) - 1u;   // n <= 2^25 keys (layout.h RTW_MAX_PRIMS)

__device__ __forceinline__ uint32_t digit_of(unsigned long long key, int shift) { return (uint32_t)(key >> shift) & 0xffu; }

__global__ void __launch_bounds__(kSortThreads) k_sort_hist_all(const unsigned long long* __restrict__ keys, int n,
                                                                 uint32_t* __restrict__ ghist) {
    __shared__ uint32_t hist[kPasses][256];
    for (int i = threadIdx.x; i < kPasses * 256; i += kSortThreads) (&hist[0][0])[i] = 0;
    __syncthreads();
    const int tiles = (n + kTile - 1) / kTile;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int base = tile * kTile;
#pragma unroll 4
        for (int r = 0; r < kItems; r++) {
            const int i = base + r * kSortThreads + threadIdx.x;
            if (i < n) {
                const unsigned long long k = keys[i];
#pragma unroll
                for (int p = 0; p < kPasses; p++) atomicAdd(&hist[p][digit_of(k, 8 * p)], 1u);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kPasses * 256; i += kSortThreads) {
        const uint32_t c = (&hist[0][0])[i];
        if (c) atomicAdd(ghist + i, c);
    }
}

__global__ void __launch_bounds__(kSortThreads) k_sort_onesweep(const unsigned long long* __restrict__ keys_in,
                                                                 const uint32_t* __restrict__ vals_in,
                                                                 unsigned long long* __restrict__ keys_out,
                                                                 uint32_t* __restrict__ vals_out, int n, int pass,
                                                                 uint32_t* __restrict__ scratch) {
    __shared__ uint32_t warp_hist[kSortWarps][256];
    __shared__ uint32_t s_scan[kSortWarps];
    __shared__ int s_tile;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int shift = 8 * pass;
    const uint32_t tag_agg = (uint32_t)(2 * pass + 1) << 27, tag_pre = (uint32_t)(2 * pass + 2) << 27;
    volatile uint32_t* status = scratch + kOffStatus;

    if (threadIdx.x == 0) s_tile = (int)atomicAdd(scratch + kOffTicket + pass, 1u);
    for (int i = threadIdx.x; i < kSortWarps * 256; i += kSortThreads) (&warp_hist[0][0])[i] = 0;
    __syncthreads();
    const int tile = s_tile;

    // global base of digit d = exclusive scan of the pass's histogram (256 values, one per thread)
    uint32_t gbase;
    {
        const uint32_t c = scratch[pass * 256 + threadIdx.x];
        uint32_t inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += v;
        }
        if (lane == 31) s_scan[warp] = inc;
        __syncthreads();
        uint32_t before = 0;
#pragma unroll
        for (int w = 0; w < kSortWarps; w++) before += (w < warp) ? s_scan[w] : 0u;
        gbase = before + inc - c;
    }

    // rank the tile's keys: a warp owns a contiguous run of kItems * 32 keys, in index order
    const int base = tile * kTile + warp * (kItems * 32);
    unsigned long long key[kItems];
    uint32_t rank[kItems];
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < kItems; r++) {
        const int i = base + r * 32 + lane;
        key[r] = i < n ? keys_in[i] : 0ull;
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
    {   // thread d: scan digit d over the warps, publish the tile's count, look back for the earlier tiles
        const int d = threadIdx.x;
        uint32_t count = 0;
#pragma unroll
        for (int w = 0; w < kSortWarps; w++) {
            const uint32_t c = warp_hist[w][d];
            warp_hist[w][d] = count;
            count += c;
        }
        uint32_t excl = 0;
        if (tile == 0) {
            status[d] = tag_pre | count;
        } else {
            status[(size_t)tile * 256 + d] = tag_agg | count;
            int t = tile - 1;
            for (;;) {
                const uint32_t s = status[(size_t)t * 256 + d];
                const uint32_t tag = s & ~kCountMask;
                if (tag == tag_pre) { excl += s & kCountMask; break; }
                if (tag == tag_agg) { excl += s & kCountMask; t--; }   // tile 0 always publishes a prefix: t stays >= 0
            }
            status[(size_t)tile * 256 + d] = tag_pre | (excl + count);
        }
        const uint32_t off = gbase + excl;
#pragma unroll
        for (int w = 0; w < kSortWarps; w++) warp_hist[w][d] += off;
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

}  // namespace

int sort_num_tiles(int n) { return (n + kTile - 1) / kTile; }
size_t sort_scratch_words(int n) { return (size_t)kOffStatus + (size_t)256 * (size_t)(sort_num_tiles(n) > 0 ? sort_num_tiles(n) : 1); }

cudaError_t sort_pairs(cudaStream_t s, unsigned long long* keys[2], uint32_t* vals[2], int n, uint32_t* scratch,
                       uint64_t* launches) {
    if (n <= 0) return cudaSuccess;
    const int tiles = sort_num_tiles(n);
    cudaError_t e = cudaMemsetAsync(scratch, 0, sizeof(uint32_t) * sort_scratch_words(n), s);
    if (e != cudaSuccess) return e;
    const int hist_ctas = tiles < 296 ? tiles : 296;
    k_sort_hist_all<<<hist_ctas, kSortThreads, 0, s>>>(keys[0], n, scratch);
    int cur = 0;
    for (int pass = 0; pass < kPasses; pass++) {
        k_sort_onesweep<<<tiles, kSortThreads, 0, s>>>(keys[cur], vals[cur], keys[cur ^ 1], vals[cur ^ 1], n, pass, scratch);
        cur ^= 1;
    }
    if (launches) *launches += 1 + kPasses;
    // 8 passes: the result is back in buffer 0
    return cudaGetLastError();
}

}  // namespace rtw
