// stream.cu — the STREAMING wavefront kernels (sm_100a).
//
// One iteration (replaces one trip of the `rayColor` while-loop of raygen/raygen.cu:36-84 for
// every path in flight):
//   k_stream_fill    start as many new camera paths as the pool has room for: tea<64> seed, jitter,
//                    camera ray (raygen.cu:129-135, shaders/camera.cu:11-19), closest hit of that ray,
//                    append the path to the queue of the material class it hit
//   k_stream_shade   the four scattering queues in one launch, one BSDF per 256-entry chunk: BSDF
//                    sample, NEE sample and ITS shadow ray, Russian roulette, the closest hit of the
//                    next ray, append to the queue of the class that ray hit
//   k_stream_advance retire the ordinals handed out, ray statistics, size of the next fill (1 warp)
// A path's state (device/stream.h, 80 bytes) is read once and written once per bounce, as consecutive
// float4 per lane; no slot pool, no index queues, no gathers.  Terminal vertices (miss, light) are
// shaded by the thread that traced the ray; finished samples go straight to the accumulator with one
// red.global.add.v4.f32.  Both heavy kernels mix the issue-bound walk over the scene with the
// latency-bound state stream, so warps of one hide the other.
#include "internal.h"
#include "../device/stream.h"

namespace rtw {

namespace {

constexpr int kBlock = 256;

// How a thread finds its hits.  `active` = this lane really has a ray (the others walk an idle ray
// when the walk is warp-uniform, or skip it when it is not).
struct TravStreamSmall {      // scene of <= RTW_SMALL_MAX leaves, by value in the constant bank: warp-uniform walk
    const SmallScene& ss;
    __device__ __forceinline__ bool closest(const RayW& r, bool, float& t, uint32_t& leaf) const { return small_closest(ss, r, t, leaf); }
    __device__ __forceinline__ bool any(const RayW& r, bool) const { return small_any(ss, r); }
};
struct TravStreamBvh {        // LBVH in global memory, top levels in shared memory: per-lane while-while walk
    const DScene& sc;
    const float4* top;
    __device__ __forceinline__ bool closest(const RayW& r, bool active, float& t, uint32_t& leaf) const {
        t = 0.f;
        leaf = RTW_REF_NONE;
        return active && traverse_closest<MemGlobal, false>(sc, top, r, t, leaf, nullptr);
    }
    __device__ __forceinline__ bool any(const RayW& r, bool active) const {
        return active && traverse_any<MemGlobal, false>(sc, top, r, nullptr);
    }
};

// After a closest-hit walk: terminal vertices are shaded and flushed here, the other paths are
// appended to the queue of their class in buffer `buf` — ONE atomic round trip per warp for all
// classes (the lanes of a class elect a leader, the leaders issue their atomicAdd together).
__device__ __forceinline__ void append_traced(const DScene& sc, const DWave& w, int buf, bool active, PathV& p, bool found,
                                              float t, uint32_t leaf) {
    const uint32_t lane = threadIdx.x & 31u;
    int cls = -1;
    if (active) cls = classify_or_finish<MemGlobal>(sc, w, p, found, t, leaf);
    __syncwarp();
    const uint32_t peers = __match_any_sync(0xffffffffu, cls);
    const int leader = __ffs(peers) - 1;
    uint32_t base = 0;
    if (cls >= 0 && (int)lane == leader) base = atomicAdd(w.counters + RTW_C_SQ0 + buf * RTW_SQ_CLASSES + cls, (uint32_t)__popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (cls >= 0) sq_store(w, buf, cls, base + __popc(peers & ((1u << lane) - 1u)), p);
}

template <class Trav>
__device__ __forceinline__ void stream_fill(const DScene& sc, const DWave& w, const Trav& trav) {
    const uint32_t n = w.counters[RTW_C_NEW];
    const unsigned long long next = *reinterpret_cast<const unsigned long long*>(w.counters + RTW_C_NEXT_LO);
    const uint32_t tiles = (n + kBlock - 1) / kBlock;
    for (uint32_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const uint32_t i = tile * kBlock + threadIdx.x;
        const bool valid = i < n;
        PathV p{};
        RayW r = idle_ray(1e-6f, 1.e27f);
        if (valid) {
            p = path_of(make_path(sc, w, next + (unsigned long long)i));
            r = closest_ray(p);
        }
        __syncwarp();
        float t;
        uint32_t leaf;
        const bool found = trav.closest(r, valid, t, leaf);
        append_traced(sc, w, w.parity, valid, p, found, t, leaf);
    }
}

template <class Trav>
__device__ __forceinline__ void stream_shade(const DScene& sc, const DWave& w, const Trav& trav, uint32_t class_mask, bool has_shadow) {
    __shared__ uint32_t s_n[RTW_SQ_CLASSES], s_first[RTW_SQ_CLASSES + 1];
    const int buf = w.parity;
    if (threadIdx.x == 0) {
        uint32_t f = 0;
        for (int c = 0; c < RTW_SQ_CLASSES; c++) {
            const uint32_t n = (class_mask >> c) & 1u ? w.counters[RTW_C_SQ0 + buf * RTW_SQ_CLASSES + c] : 0u;
            s_n[c] = n;
            s_first[c] = f;
            f += (n + kBlock - 1) / kBlock;
        }
        s_first[RTW_SQ_CLASSES] = f;
    }
    __syncthreads();
    const uint32_t chunks = s_first[RTW_SQ_CLASSES];
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t n_traced = 0, n_shadow = 0;   // this warp's rays (same value in every lane)
    for (uint32_t v = blockIdx.x; v < chunks; v += gridDim.x) {
        int c = 0;
#pragma unroll
        for (int k = 1; k < RTW_SQ_CLASSES; k++) c += (v >= s_first[k]) ? 1 : 0;
        const uint32_t i = (v - s_first[c]) * kBlock + threadIdx.x;
        const bool valid = i < s_n[c];
        PathV p{};
        ShadowOut sh;
        sh.emit = false;
        bool alive = false;
        if (valid) {
            p = sq_load(w, buf, c, i);
            alive = shade_core<MemGlobal>(sc, w.max_depth, c, p, sh);
        }
        __syncwarp();
        // NEE shadow ray (traceOcclusion, shaders/closehit.cu:16-42,97-113: tmin 500e-7, ray time 0)
        if (has_shadow && c == RTW_Q_LAMBERTIAN) {
            const uint32_t m_shadow = __ballot_sync(0xffffffffu, sh.emit);
            if (m_shadow) {
                RayW r = idle_ray(500 * 1.0e-7f, 0.f);
                if (sh.emit) {
                    r.o = sh.o; r.d = sh.d; r.tmax = sh.tmax;
                    r.gather = sc.has_motion ? p.gather : 0.f;
                    r.seed = sh.seed;
                }
                const bool occluded = trav.any(r, sh.emit);
                if (sh.emit && !occluded) {
                    p.rad.x += sh.contrib.x; p.rad.y += sh.contrib.y; p.rad.z += sh.contrib.z;
                }
                n_shadow += (uint32_t)__popc(m_shadow);
            }
        }
        if (valid && !alive) path_flush(w, p);
        const uint32_t m_alive = __ballot_sync(0xffffffffu, alive);
        if (m_alive) {
            RayW r = idle_ray(1e-6f, 1.e27f);
            if (alive) r = closest_ray(p);
            float t;
            uint32_t leaf;
            const bool found = trav.closest(r, alive, t, leaf);
            append_traced(sc, w, buf ^ 1, alive, p, found, t, leaf);
            n_traced += (uint32_t)__popc(m_alive);
        }
    }
    if (lane == 0 && n_traced) atomicAdd(w.counters + RTW_C_TRACED, n_traced);
    if (lane == 1 && n_shadow) atomicAdd(w.counters + RTW_C_SHADOW, n_shadow);
}

__global__ void __launch_bounds__(kBlock) k_stream_fill_small(const DScene sc, const DWave w, const __grid_constant__ SmallScene ss) {
    stream_fill(sc, w, TravStreamSmall{ss});
}
__global__ void __launch_bounds__(kBlock) k_stream_fill_bvh(const DScene sc, const DWave w) {
    extern __shared__ float4 s_top[];
    const int n4 = 4 * sc.n_top;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) s_top[i] = __ldg(sc.top_nodes + i);
    __syncthreads();
    stream_fill(sc, w, TravStreamBvh{sc, s_top});
}
template <int kCtas>
__global__ void __launch_bounds__(kBlock, kCtas) k_stream_shade_small(const DScene sc, const DWave w, const __grid_constant__ SmallScene ss,
                                                                      uint32_t class_mask, int has_shadow) {
    stream_shade(sc, w, TravStreamSmall{ss}, class_mask, has_shadow != 0);
}
template <int kCtas>
__global__ void __launch_bounds__(kBlock, kCtas) k_stream_shade_bvh(const DScene sc, const DWave w, uint32_t class_mask, int has_shadow) {
    extern __shared__ float4 s_top[];
    const int n4 = 4 * sc.n_top;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) s_top[i] = __ldg(sc.top_nodes + i);
    __syncthreads();
    stream_shade(sc, w, TravStreamBvh{sc, s_top}, class_mask, has_shadow != 0);
}

// After the shading kernel of iteration `parity`: buffer `parity` has been consumed, buffer
// `parity ^ 1` holds the paths that go on.
__global__ void k_stream_advance(const DWave w) {
    if (threadIdx.x == 0) {
        uint32_t* c = w.counters;
        unsigned long long* next = reinterpret_cast<unsigned long long*>(c + RTW_C_NEXT_LO);
        unsigned long long* rays = reinterpret_cast<unsigned long long*>(c + RTW_C_RAYS_LO);
        unsigned long long* shrays = reinterpret_cast<unsigned long long*>(c + RTW_C_SHRAYS_LO);
        const uint32_t started = c[RTW_C_NEW];
        *next += started;
        *rays += (unsigned long long)started + (unsigned long long)c[RTW_C_TRACED];
        *shrays += (unsigned long long)c[RTW_C_SHADOW];
        uint32_t alive = 0;
        for (int k = 0; k < RTW_SQ_CLASSES; k++) {
            alive += c[RTW_C_SQ0 + (w.parity ^ 1) * RTW_SQ_CLASSES + k];
            c[RTW_C_SQ0 + w.parity * RTW_SQ_CLASSES + k] = 0u;
        }
        const unsigned long long left = *next < w.total ? w.total - *next : 0ull;
        const unsigned long long room = (unsigned long long)((uint32_t)w.P - alive);
        const uint32_t fill = (uint32_t)(left < room ? left : room);
        c[RTW_C_NEW] = fill;
        c[RTW_C_ALIVE] = alive + fill;   // 0 ⇒ nothing in flight, nothing left to start: the host stops iterating
        c[RTW_C_TRACED] = 0u;
        c[RTW_C_SHADOW] = 0u;
    }
}

__global__ void k_stream_init(const DWave w) {
    if (threadIdx.x < RTW_C_WORDS) w.counters[threadIdx.x] = 0u;
    __syncwarp();
    if (threadIdx.x == 0) {
        const unsigned long long room = (unsigned long long)w.P;
        const uint32_t fill = (uint32_t)(w.total < room ? w.total : room);
        w.counters[RTW_C_NEW] = fill;
        w.counters[RTW_C_ALIVE] = fill;
    }
}

// ---------------------------------------------------------------------------
// Drain tail of a SLOT-POOL frame (kernels/wavefront.cu) in one launch.
//
// When no sample is left to start, the population of a pool decays by ~30 % per iteration and the last
// ~30 iterations each cost four launches for a few thousand rays: launch latency, not work (0.6-0.9 ms per
// frame, 5 % of the 17 ms frame one GPU of eight renders).  Once fewer than `tail_paths` paths are alive the
// host stops iterating and launches k_tail_flush (the slots that finished in the last iteration go to the
// accumulator, as k_regen would do) and k_tail_*: one THREAD per surviving path runs it to its end — trace,
// terminal vertex or BSDF + NEE + shadow ray + Russian roulette, next ray — with the per-thread bodies of the
// streaming wavefront (device/stream.h), i.e. the same statements in the same order as the staged kernels.
__device__ __forceinline__ PathV path_of_slot(const DWave& w, uint32_t slot) {
    const RayRec ray = w.rec[slot].ray;
    const PathRec st = w.rec[slot].st;
    const float4 ra = w.rad[slot];
    PathV p;
    p.o = f3(ray.o); p.time = ray.o.w;
    p.d = f3(ray.d); p.gather = ray.d.w;
    p.thr = f3(st.thr); p.seed = f2u(st.thr.w);
    p.depth = misc_depth(st.misc.x); p.lseed = st.misc.y;
    p.t = 0.f; p.leaf = RTW_REF_NONE;
    p.rad = f3(ra); p.pixel = f2u(ra.w);
    return p;
}

__global__ void __launch_bounds__(kBlock) k_tail_flush(const DWave w) {
    const uint32_t n = w.counters[RTW_C_QUEUE0 + RTW_Q_REGEN];
    const uint32_t* q = w.queue + (size_t)RTW_Q_REGEN * w.P;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t slot = q[i];
        logic_flush(w, w.rad[slot]);
        retire_slot(w, slot);      // k_tail_* may find its paths by scanning the pool: this slot holds none any more
    }
}

template <class Trav>
__device__ __forceinline__ void tail_run(const DScene& sc, const DWave& w, const Trav& trav, bool has_shadow) {
    // the paths: the list the last shading launch kept (sparse mode), or every slot that holds a ray
    const bool listed = w.trace_use != 0;
    const uint32_t n = listed ? w.counters[RTW_C_TRACE0 + w.parity] : (uint32_t)w.P;
    const uint32_t* list = w.queue + (size_t)RTW_Q_TRACE * w.P;
    const uint32_t tiles = (n + kBlock - 1) / kBlock;
    uint32_t n_closest = 0, n_shadow = 0;
    for (uint32_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const uint32_t i = tile * kBlock + threadIdx.x;
        bool alive = i < n;
        PathV p{};
        if (alive) {
            const uint32_t slot = listed ? list[i] : i;
            alive = f2u(w.rec[slot].ray.o.w) != RTW_RAY_NONE;
            if (alive) p = path_of_slot(w, slot);
        }
        while (__any_sync(0xffffffffu, alive)) {
            RayW r = idle_ray(1e-6f, 0.f);
            if (alive) r = closest_ray(p);
            float t;
            uint32_t leaf;
            const bool found = trav.closest(r, alive, t, leaf);
            int cls = -1;
            if (alive) {
                n_closest++;
                cls = classify_or_finish<MemGlobal>(sc, w, p, found, t, leaf);   // terminal vertex: shaded and flushed
                if (cls < 0) alive = false;
            }
            ShadowOut sh;
            sh.emit = false;
            bool goes_on = false;
            if (alive) goes_on = shade_core<MemGlobal>(sc, w.max_depth, cls, p, sh);
            __syncwarp();
            if (has_shadow && __any_sync(0xffffffffu, sh.emit)) {
                RayW q = idle_ray(500 * 1.0e-7f, 0.f);
                if (sh.emit) {
                    q.o = sh.o; q.d = sh.d; q.tmax = sh.tmax;
                    q.gather = sc.has_motion ? p.gather : 0.f;
                    q.seed = sh.seed;
                    n_shadow++;
                }
                const bool occluded = trav.any(q, sh.emit);
                if (sh.emit && !occluded) {
                    p.rad.x += sh.contrib.x; p.rad.y += sh.contrib.y; p.rad.z += sh.contrib.z;
                }
            }
            if (alive && !goes_on) {
                path_flush(w, p);
                alive = false;
            }
        }
    }
    n_closest = __reduce_add_sync(0xffffffffu, n_closest);
    n_shadow = __reduce_add_sync(0xffffffffu, n_shadow);
    if ((threadIdx.x & 31u) == 0) {
        if (n_closest) atomicAdd(reinterpret_cast<unsigned long long*>(w.counters + RTW_C_RAYS_LO), (unsigned long long)n_closest);
        if (n_shadow) atomicAdd(reinterpret_cast<unsigned long long*>(w.counters + RTW_C_SHRAYS_LO), (unsigned long long)n_shadow);
    }
}

__global__ void __launch_bounds__(kBlock) k_tail_small(const DScene sc, const DWave w, const __grid_constant__ SmallScene ss, int has_shadow) {
    tail_run(sc, w, TravStreamSmall{ss}, has_shadow != 0);
}
__global__ void __launch_bounds__(kBlock) k_tail_bvh(const DScene sc, const DWave w, int has_shadow) {
    extern __shared__ float4 s_top[];
    const int n4 = 4 * sc.n_top;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) s_top[i] = __ldg(sc.top_nodes + i);
    __syncthreads();
    tail_run(sc, w, TravStreamBvh{sc, s_top}, has_shadow != 0);
}

template <class K>
int resident_grid(K kernel, size_t smem, int sms) {
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kBlock, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
    return sms * per_sm;
}

}  // namespace

cudaError_t stream_configure(const DScene& sc, LaunchCfg* cfg) {
    (void)sc;
    int dev = 0, sms = 148;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cfg->small) {
        cfg->grid_fill = resident_grid(k_stream_fill_small, 0, sms);
        cfg->grid_stream = cfg->shade_ctas == 2 ? resident_grid(k_stream_shade_small<2>, 0, sms)
                         : cfg->shade_ctas == 4 ? resident_grid(k_stream_shade_small<4>, 0, sms)
                                                : resident_grid(k_stream_shade_small<3>, 0, sms);
    } else {
        if (cfg->smem_top > 48 * 1024) {
            e = cudaFuncSetAttribute(k_stream_fill_bvh, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg->smem_top);
            if (e != cudaSuccess) return e;
            e = cudaFuncSetAttribute(k_tail_bvh, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg->smem_top);
            if (e != cudaSuccess) return e;
            e = cudaFuncSetAttribute(k_stream_shade_bvh<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg->smem_top);
            if (e != cudaSuccess) return e;
            e = cudaFuncSetAttribute(k_stream_shade_bvh<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg->smem_top);
            if (e != cudaSuccess) return e;
            e = cudaFuncSetAttribute(k_stream_shade_bvh<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg->smem_top);
            if (e != cudaSuccess) return e;
        }
        cfg->grid_fill = resident_grid(k_stream_fill_bvh, cfg->smem_top, sms);
        cfg->grid_stream = cfg->shade_ctas == 2 ? resident_grid(k_stream_shade_bvh<2>, cfg->smem_top, sms)
                         : cfg->shade_ctas == 4 ? resident_grid(k_stream_shade_bvh<4>, cfg->smem_top, sms)
                                                : resident_grid(k_stream_shade_bvh<3>, cfg->smem_top, sms);
    }
    return cudaSuccess;
}

cudaError_t launch_tail(cudaStream_t s, const DScene& sc, const DWave& w, const LaunchCfg& cfg, uint64_t* launches) {
    k_tail_flush<<<148, kBlock, 0, s>>>(w);
    const int grid = cfg.grid_fill > 0 ? cfg.grid_fill : 148;
    if (cfg.small) k_tail_small<<<grid, kBlock, 0, s>>>(sc, w, *cfg.small, cfg.has_shadow ? 1 : 0);
    else k_tail_bvh<<<grid, kBlock, cfg.smem_top, s>>>(sc, w, cfg.has_shadow ? 1 : 0);
    if (launches) *launches += 2;
    return cudaGetLastError();
}

cudaError_t stream_init(cudaStream_t s, const DWave& w, uint64_t* launches) {
    k_stream_init<<<1, 32, 0, s>>>(w);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

cudaError_t stream_iteration(cudaStream_t s, const DScene& sc, const DWave& w, const LaunchCfg& cfg, uint64_t* launches,
                             cudaEvent_t* ev) {
    const uint32_t classes = cfg.class_mask;
    const int shadow = cfg.has_shadow ? 1 : 0;
    if (ev) cudaEventRecord(ev[0], s);
    if (cfg.small) k_stream_fill_small<<<cfg.grid_fill, kBlock, 0, s>>>(sc, w, *cfg.small);
    else k_stream_fill_bvh<<<cfg.grid_fill, kBlock, cfg.smem_top, s>>>(sc, w);
    if (cfg.mid_event) cudaEventRecord(cfg.mid_event, s);
    if (ev) { cudaEventRecord(ev[1], s); cudaEventRecord(ev[2], s); }
    if (cfg.small) {
        if (cfg.shade_ctas == 2) k_stream_shade_small<2><<<cfg.grid_stream, kBlock, 0, s>>>(sc, w, *cfg.small, classes, shadow);
        else if (cfg.shade_ctas == 4) k_stream_shade_small<4><<<cfg.grid_stream, kBlock, 0, s>>>(sc, w, *cfg.small, classes, shadow);
        else k_stream_shade_small<3><<<cfg.grid_stream, kBlock, 0, s>>>(sc, w, *cfg.small, classes, shadow);
    } else {
        if (cfg.shade_ctas == 2) k_stream_shade_bvh<2><<<cfg.grid_stream, kBlock, cfg.smem_top, s>>>(sc, w, classes, shadow);
        else if (cfg.shade_ctas == 4) k_stream_shade_bvh<4><<<cfg.grid_stream, kBlock, cfg.smem_top, s>>>(sc, w, classes, shadow);
        else k_stream_shade_bvh<3><<<cfg.grid_stream, kBlock, cfg.smem_top, s>>>(sc, w, classes, shadow);
    }
    if (ev) cudaEventRecord(ev[3], s);
    k_stream_advance<<<1, 32, 0, s>>>(w);
    if (ev) cudaEventRecord(ev[4], s);
    if (launches) *launches += 3;
    return cudaGetLastError();
}

}  // namespace rtw
