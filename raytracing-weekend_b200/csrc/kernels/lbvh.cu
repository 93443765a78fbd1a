// lbvh.cu — GPU LBVH build (replaces optixAccelBuild).
//
//   k_prim_aabb   world AABB per primitive + scene bounds (warp shuffle + atomics)
//   k_morton      63-bit Morton code of the AABB centroid
//   sort_pairs    stable radix sort (kernels/sort.cu)
//   k_pack_leaves leaf-ordered primitive records (32 B), leaf boxes, ids
//   k_karras      Karras 2012 hierarchy, one thread per internal node
//   k_refit       bottom-up AABB refit with atomic arrival flags
//   k_build_top   breadth-first copy of the top nodes for shared-memory staging
// All kernels are HBM-bound streaming passes; bytes per primitive are listed in
// DESIGN.md ("LBVH build").
#include "internal.h"
#include "../device/lbvh.h"

namespace rtw {

namespace {

constexpr int kThreads = 256;

// order-preserving float → uint map for atomicMin/atomicMax
__device__ __forceinline__ uint32_t enc_f(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float dec_f(uint32_t u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// bounds[0..5] scene box (encoded min xyz, max xyz), [6] n_top, [8..13] bounds of the box centroids,
// [14] primitives flagged oversize (lbvh.h); 16 words
__global__ void k_bounds_init(uint32_t* bounds) {
    const int t = threadIdx.x & 7;
    if (threadIdx.x < 16) bounds[threadIdx.x] = t < 3 ? 0xffffffffu : 0u;   // min words start at the top, everything else at 0
}

__global__ void __launch_bounds__(kThreads) k_prim_aabb(const rtw_prim* __restrict__ prims,
                                                         const rtw_transform* __restrict__ xforms, int n,
                                                         float4* __restrict__ box_mn, float4* __restrict__ box_mx,
                                                         uint32_t* __restrict__ bounds) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    // v[0..2] min, v[3..5] max of the boxes; v[6..8] min, v[9..11] max of the box centroids
    float v[12];
#pragma unroll
    for (int k = 0; k < 12; k++) v[k] = (k % 6) < 3 ? RTW_FLT_MAX : -RTW_FLT_MAX;
    if (i < n) {
        const Box3 b = prim_world_aabb(prims[i], xforms);
        box_mn[i] = make_float4(b.mn.x, b.mn.y, b.mn.z, 0.f);
        box_mx[i] = make_float4(b.mx.x, b.mx.y, b.mx.z, 0.f);
        const float3 c = box_centroid(b);
        v[0] = b.mn.x; v[1] = b.mn.y; v[2] = b.mn.z; v[3] = b.mx.x; v[4] = b.mx.y; v[5] = b.mx.z;
        v[6] = c.x; v[7] = c.y; v[8] = c.z; v[9] = c.x; v[10] = c.y; v[11] = c.z;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int k = 0; k < 12; k++) {
            const float u = __shfl_xor_sync(0xffffffffu, v[k], o);
            v[k] = (k % 6) < 3 ? fminf(v[k], u) : fmaxf(v[k], u);
        }
    }
    // one set of atomics per CTA, not per warp: at 2^20 primitives 32768 warps x 6 atomics on the same six words
    // were 140 us of serialised read-modify-writes (profiles/r02/build_r02f.md)
    __shared__ float s_b[kThreads / 32][12];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 12; k++) s_b[warp][k] = v[k];
    }
    __syncthreads();
    if (threadIdx.x < 12) {
        const int k = threadIdx.x;
        const bool is_min = (k % 6) < 3;
        float r = s_b[0][k];
#pragma unroll
        for (int w = 1; w < kThreads / 32; w++) r = is_min ? fminf(r, s_b[w][k]) : fmaxf(r, s_b[w][k]);
        uint32_t* word = bounds + (k < 6 ? k : k + 2);
        if (is_min) atomicMin(word, enc_f(r));
        else atomicMax(word, enc_f(r));
    }
}

__global__ void __launch_bounds__(kThreads) k_morton(const float4* __restrict__ box_mn, const float4* __restrict__ box_mx,
                                                      int n, uint32_t* bounds,
                                                      unsigned long long* __restrict__ keys, uint32_t* __restrict__ vals) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float3 lo = f3(dec_f(bounds[0]), dec_f(bounds[1]), dec_f(bounds[2]));
    const float3 hi = f3(dec_f(bounds[3]), dec_f(bounds[4]), dec_f(bounds[5]));
    const float3 clo = f3(dec_f(bounds[8]), dec_f(bounds[9]), dec_f(bounds[10]));
    const float3 chi = f3(dec_f(bounds[11]), dec_f(bounds[12]), dec_f(bounds[13]));
    Box3 b;
    b.mn = f3(box_mn[i]);
    b.mx = f3(box_mx[i]);
    unsigned long long key = morton63(b, lo, hi - lo);
    if (box_is_oversize(b, clo, chi)) {   // sorts behind every other key (lbvh.h "oversize primitives")
        key |= RTW_KEY_OVERSIZE;
        atomicAdd(bounds + 14, 1u);
    }
    keys[i] = key;
    vals[i] = (uint32_t)i;
}

__global__ void __launch_bounds__(kThreads) k_pack_leaves(const rtw_prim* __restrict__ prims,
                                                           const rtw_material* __restrict__ mats, int n,
                                                           const uint32_t* __restrict__ vals,
                                                           const float4* __restrict__ box_mn, const float4* __restrict__ box_mx,
                                                           float4* __restrict__ leaf_mn, float4* __restrict__ leaf_mx,
                                                           float4* __restrict__ packed, float2* __restrict__ ext,
                                                           int* __restrict__ leaf_prim, int* __restrict__ leaf_mat) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const int pi = (int)vals[j];
    const rtw_prim p = prims[pi];
    float4 a, b;
    float2 e;
    pack_prim(p, mats[p.material].type, a, b, e);
    packed[2 * j] = a;
    packed[2 * j + 1] = b;
    ext[j] = e;
    leaf_prim[j] = pi;
    leaf_mat[j] = p.material;
    leaf_mn[j] = box_mn[pi];
    leaf_mx[j] = box_mx[pi];
}

__global__ void __launch_bounds__(kThreads) k_karras(const unsigned long long* __restrict__ keys, int n_all,
                                                      const uint32_t* __restrict__ bounds,
                                                      float4* __restrict__ nodes, int* __restrict__ parent,
                                                      int* __restrict__ leaf_parent, int* __restrict__ flags,
                                                      int* __restrict__ top_index, int leaf_size) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = tree_leaf_count(n_all, (int)bounds[14]);   // the hierarchy covers the leaves in front of the oversize list
    if (i >= n - 1) return;
    int left, right, lo, gamma, hi;
    karras_node(keys, n, i, left, right, lo, gamma, hi);
    if (left < 0) leaf_parent[~left] = i; else parent[left] = i;
    if (right < 0) leaf_parent[~right] = i; else parent[right] = i;
    if (i == 0) parent[0] = -1;
    flags[i] = 0;
    top_index[i] = -1;
    // children now, boxes at refit.  n3.xy = traversal refs (subtrees of <= leaf_size
    // primitives collapsed into one leaf range), n3.zw = the Karras hierarchy itself
    // (refit, topology download).
    nodes[4 * i + 3] = make_float4(__uint_as_float(traversal_ref(left, lo, gamma, leaf_size)),
                                   __uint_as_float(traversal_ref(right, gamma + 1, hi, leaf_size)),
                                   __uint_as_float(encode_ref(left)), __uint_as_float(encode_ref(right)));
}

// One-leaf scene: node 0 = (leaf 0, unreachable point box at +FLT_MAX).  An inverted
// "empty" box would not work: the slab test symmetrises min/max per axis.
__global__ void k_single_leaf(const float4* leaf_mn, const float4* leaf_mx, float4* nodes, int* parent, int* top_index) {
    const float4 mn = leaf_mn[0], mx = leaf_mx[0];
    nodes[0] = make_float4(mn.x, mn.y, mn.z, mx.x);
    nodes[1] = make_float4(mx.y, mx.z, RTW_FLT_MAX, RTW_FLT_MAX);
    nodes[2] = make_float4(RTW_FLT_MAX, RTW_FLT_MAX, RTW_FLT_MAX, RTW_FLT_MAX);
    nodes[3] = make_float4(__uint_as_float(RTW_REF_LEAF | 0u), __uint_as_float(RTW_REF_LEAF | 0u),
                           __uint_as_float(RTW_REF_LEAF | 0u), __uint_as_float(RTW_REF_LEAF | 0u));
    parent[0] = -1;
    top_index[0] = -1;
}

__device__ __forceinline__ void child_box(uint32_t ref, const float4* leaf_mn, const float4* leaf_mx,
                                          const float4* nodes, float3& mn, float3& mx) {
    if (ref & RTW_REF_LEAF) {
        const uint32_t l = ref & RTW_REF_MASK;
        mn = f3(leaf_mn[l]);
        mx = f3(leaf_mx[l]);
    } else {
        // written by another thread before its atomic arrival: bypass L1
        const float4 n0 = __ldcg(nodes + 4 * ref), n1 = __ldcg(nodes + 4 * ref + 1), n2 = __ldcg(nodes + 4 * ref + 2);
        mn = vmin(f3(n0.x, n0.y, n0.z), f3(n1.z, n1.w, n2.x));
        mx = vmax(f3(n0.w, n1.x, n1.y), f3(n2.y, n2.z, n2.w));
    }
}

__global__ void __launch_bounds__(kThreads) k_refit(int n_all, const uint32_t* __restrict__ bounds,
                                                     const int* __restrict__ leaf_parent, const int* __restrict__ parent,
                                                     int* flags, const float4* __restrict__ leaf_mn,
                                                     const float4* __restrict__ leaf_mx, float4* nodes) {
    const int leaf = blockIdx.x * blockDim.x + threadIdx.x;
    if (leaf >= tree_leaf_count(n_all, (int)bounds[14])) return;
    int cur = leaf_parent[leaf];
    while (cur >= 0) {
        if (atomicAdd(flags + cur, 1) == 0) return;  // first arrival stops, second computes
        __threadfence();
        const float4 n3 = nodes[4 * cur + 3];
        float3 lmn, lmx, rmn, rmx;
        child_box(__float_as_uint(n3.z), leaf_mn, leaf_mx, nodes, lmn, lmx);
        child_box(__float_as_uint(n3.w), leaf_mn, leaf_mx, nodes, rmn, rmx);
        nodes[4 * cur + 0] = make_float4(lmn.x, lmn.y, lmn.z, lmx.x);
        nodes[4 * cur + 1] = make_float4(lmx.y, lmx.z, rmn.x, rmn.y);
        nodes[4 * cur + 2] = make_float4(rmn.z, rmx.x, rmx.y, rmx.z);
        __threadfence();
        cur = parent[cur];
    }
}

// Breadth-first order of the first max_top internal nodes; children that are
// themselves staged are re-pointed at the shared-memory table.
__global__ void __launch_bounds__(kThreads) k_build_top(const float4* __restrict__ nodes, int n_internal, int max_top,
                                                         int* __restrict__ top_index, float4* __restrict__ top_nodes,
                                                         uint32_t* __restrict__ n_top_out) {
    extern __shared__ int queue[];
    __shared__ int s_count;
    if (threadIdx.x == 0) {
        int head = 0, tail = 0;
        if (max_top > 0) {
            queue[tail] = 0;
            top_index[0] = 0;
            tail = 1;
        }
        while (head < tail) {
            const int g = queue[head++];
            const float4 n3 = nodes[4 * g + 3];
            const uint32_t c[2] = {__float_as_uint(n3.x), __float_as_uint(n3.y)};
            for (int k = 0; k < 2; k++) {
                if (!(c[k] & RTW_REF_LEAF) && tail < max_top && (int)c[k] < n_internal && top_index[c[k]] < 0) {
                    top_index[c[k]] = tail;
                    queue[tail++] = (int)c[k];
                }
            }
        }
        s_count = tail;
        *n_top_out = (uint32_t)tail;
        __threadfence_block();
    }
    __syncthreads();
    const int count = s_count;
    for (int j = threadIdx.x; j < count; j += blockDim.x) {
        const int g = queue[j];
        float4 n3 = nodes[4 * g + 3];
        uint32_t cl = __float_as_uint(n3.x), cr = __float_as_uint(n3.y);
        if (!(cl & RTW_REF_LEAF) && top_index[cl] >= 0) cl = RTW_REF_TOP | (uint32_t)top_index[cl];
        if (!(cr & RTW_REF_LEAF) && top_index[cr] >= 0) cr = RTW_REF_TOP | (uint32_t)top_index[cr];
        n3.x = __uint_as_float(cl);
        n3.y = __uint_as_float(cr);
        top_nodes[4 * j + 0] = nodes[4 * g + 0];
        top_nodes[4 * j + 1] = nodes[4 * g + 1];
        top_nodes[4 * j + 2] = nodes[4 * g + 2];
        top_nodes[4 * j + 3] = n3;
    }
}

// 4-wide nodes derived from the refitted two-child nodes: one thread per internal node
__global__ void __launch_bounds__(kThreads) k_widen(const float4* __restrict__ nodes, int n_all, const uint32_t* __restrict__ bounds,
                                                     float4* __restrict__ wide) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= tree_leaf_count(n_all, (int)bounds[14]) - 1) return;
    float4 w[8];
    make_wide_node(nodes, i, w);
#pragma unroll
    for (int k = 0; k < 8; k++) wide[8 * (size_t)i + k] = w[k];
}

}  // namespace

cudaError_t lbvh_build(cudaStream_t s, const BuildBuffers& b, uint64_t* launches) {
    const int n = b.n;
    const int nb = (n + kThreads - 1) / kThreads;
    uint64_t l = 0;
    k_bounds_init<<<1, 32, 0, s>>>(b.bounds);
    k_prim_aabb<<<nb, kThreads, 0, s>>>(b.prims, b.xforms, n, b.box_mn, b.box_mx, b.bounds);
    k_morton<<<nb, kThreads, 0, s>>>(b.box_mn, b.box_mx, n, b.bounds, b.keys[0], b.vals[0]);
    l += 3;
    unsigned long long* keys[2] = {b.keys[0], b.keys[1]};
    uint32_t* vals[2] = {b.vals[0], b.vals[1]};
    cudaError_t e = sort_pairs(s, keys, vals, n, b.hist, &l);
    if (e != cudaSuccess) return e;
    k_pack_leaves<<<nb, kThreads, 0, s>>>(b.prims, b.mats_raw, n, b.vals[0], b.box_mn, b.box_mx, b.leaf_mn, b.leaf_mx,
                                          b.prims_packed, b.prim_ext, b.leaf_prim, b.leaf_mat);
    l += 1;
    int n_internal;
    if (n == 1) {
        n_internal = 1;
        k_single_leaf<<<1, 1, 0, s>>>(b.leaf_mn, b.leaf_mx, b.nodes, b.parent, b.top_index);
        l += 1;
    } else {
        n_internal = n - 1;
        const int nbi = (n_internal + kThreads - 1) / kThreads;
        k_karras<<<nbi, kThreads, 0, s>>>(b.keys[0], n, b.bounds, b.nodes, b.parent, b.leaf_parent, b.flags, b.top_index,
                                          b.leaf_size < 1 ? 1 : (b.leaf_size > RTW_MAX_LEAF_SIZE ? RTW_MAX_LEAF_SIZE : b.leaf_size));
        k_refit<<<nb, kThreads, 0, s>>>(n, b.bounds, b.leaf_parent, b.parent, b.flags, b.leaf_mn, b.leaf_mx, b.nodes);
        l += 2;
    }
    // the shared-memory top table is for the two-child walk only (the wide walk reads its 128-byte records through
    // L1): its breadth-first numbering is one thread's loop — 0.67 ms for 1024 nodes — and is skipped when unused
    // (bounds[6] = n_top stays 0 from k_bounds_init)
    if (!b.wide) {
        const int max_top = b.max_top < n_internal ? b.max_top : n_internal;
        k_build_top<<<1, kThreads, sizeof(int) * (max_top > 0 ? max_top : 1), s>>>(b.nodes, n_internal, max_top, b.top_index,
                                                                                  b.top_nodes, b.bounds + 6);
        l += 1;
    }
    if (b.wide) {
        k_widen<<<(n_internal + kThreads - 1) / kThreads, kThreads, 0, s>>>(b.nodes, n, b.bounds, b.wide);
        l += 1;
    }
    if (launches) *launches += l;
    return cudaGetLastError();
}

}  // namespace rtw
