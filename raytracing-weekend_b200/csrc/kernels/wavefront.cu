// wavefront.cu — the wavefront path-tracing kernels (sm_100a).
//
// One iteration of the pool (replaces one trip of the `rayColor` while-loop of
// raygen/raygen.cu:36-84 for every live path at once):
//   k_regen       flush the paths that finished last iteration (compact queue), start new camera paths
//   k_advance     retire the sample ordinals just handed out, clear the per-iteration counters
//   k_traverse    closest-hit LBVH traversal + per-material queue compaction
//   k_shade_all   the five material queues in one launch, one BSDF per CTA chunk: BSDF, NEE sample, Russian roulette
//   k_shadow      any-hit traversal of the NEE rays, adds unoccluded contributions
// Queue pushes are warp-aggregated: one ballot + one atomicAdd per warp and class,
// so every warp of a shading kernel runs exactly one BSDF.
// Grids are persistent-style: (SM count x resident CTAs) with grid-stride loops.
#include "internal.h"
#include "../device/wavefront.h"

namespace rtw {

namespace {

constexpr int kBlock = 256;

__device__ __forceinline__ unsigned long long atomic_add_u64(uint32_t* lo_word, unsigned long long v) {
    return atomicAdd(reinterpret_cast<unsigned long long*>(lo_word), v);
}

// Copy the breadth-first top of the LBVH into shared memory.
__device__ __forceinline__ void stage_top(const DScene& sc, float4* s_top) {
    const int n4 = 4 * sc.n_top;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) s_top[i] = __ldg(sc.top_nodes + i);
    __syncthreads();
}

// ---------------------------------------------------------------------------
// Warp-aggregated push of `slot` onto queue `cls` for the lanes where `pred` holds:
// one ballot + one atomicAdd per warp.
__device__ __forceinline__ void queue_push(const DWave& w, int cls, bool pred, uint32_t slot) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t mask = __ballot_sync(0xffffffffu, pred);
    if (!mask) return;
    const int leader = __ffs(mask) - 1;
    uint32_t base = 0;
    if ((int)lane == leader) base = atomicAdd(w.counters + RTW_C_QUEUE0 + cls, (uint32_t)__popc(mask));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (pred) w.queue[(size_t)cls * w.P + base + __popc(mask & ((1u << lane) - 1u))] = slot;
}

// Regeneration: the slots whose path finished in the previous iteration (a compact
// queue, so every lane of a warp does the same work) are flushed into the accumulator
// and restarted with the next sample ordinal (raygen.cu:129-135, :141-158).
__global__ void __launch_bounds__(kBlock) k_regen(const DScene sc, const DWave w) {
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t n = w.counters[RTW_C_QUEUE0 + RTW_Q_REGEN];
    const unsigned long long next = *reinterpret_cast<const unsigned long long*>(w.counters + RTW_C_NEXT_LO);
    const uint32_t* q = w.queue + (size_t)RTW_Q_REGEN * w.P;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const uint32_t slot = q[i];
        logic_flush(w, slot);
        // queue position i takes sample ordinal next + i: no atomic (k_advance moves `next` on by n
        // after this kernel)
        const unsigned long long ordinal = next + (unsigned long long)i;
        if (ordinal < w.total) generate_path(sc, w, slot, ordinal);
    }
}

// Between k_regen and k_traverse: the sample ordinals handed out by k_regen are retired and the
// per-iteration counters (queue lengths, shadow count, rays in flight) are cleared.
__global__ void k_advance(const DWave w) {
    if (threadIdx.x == 0) {
        unsigned long long* next = reinterpret_cast<unsigned long long*>(w.counters + RTW_C_NEXT_LO);
        *next += (unsigned long long)w.counters[RTW_C_QUEUE0 + RTW_Q_REGEN];
    }
    __syncwarp();
    if (threadIdx.x < 8) w.counters[threadIdx.x] = 0u;
}

// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) k_traverse(const DScene sc, const DWave w) {
    extern __shared__ float4 s_top[];
    stage_top(sc, s_top);
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x; slot - lane < (uint32_t)w.P; slot += stride) {
        int cls = -1;
        if (slot < (uint32_t)w.P) cls = traverse_thread<MemGlobal>(sc, s_top, w, slot);
        __syncwarp();   // rays leave the traversal loops at different times: reconverge before the warp-wide push
        // one atomic round trip per warp for ALL classes: the lanes of a class elect a leader,
        // the (up to five) leaders issue their atomicAdd in the same instruction
        const uint32_t peers = __match_any_sync(0xffffffffu, cls);
        const int leader = __ffs(peers) - 1;
        uint32_t base = 0;
        if (cls >= 0 && (int)lane == leader) base = atomicAdd(w.counters + RTW_C_QUEUE0 + cls, (uint32_t)__popc(peers));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (cls >= 0) w.queue[(size_t)cls * w.P + base + __popc(peers & ((1u << lane) - 1u))] = slot;
    }
}

// ---------------------------------------------------------------------------
// One 256-wide chunk [i0, i0 + 256) of shading queue kClass: every warp runs ONE BSDF.
template <int kClass>
__device__ __forceinline__ void shade_chunk(const DScene& sc, const DWave& w, uint32_t i0, uint32_t n) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t i = i0 + threadIdx.x;
    const uint32_t* q = w.queue + (size_t)kClass * w.P;
    ShadowOut sh;
    sh.emit = false;
    bool finished = false;
    uint32_t slot = 0;
    if (i < n) {
        slot = q[i];
        finished = shade_thread<kClass, MemGlobal>(sc, w, slot, sh);
    }
    __syncwarp();
    queue_push(w, RTW_Q_REGEN, finished, slot);
    if (kClass == RTW_Q_LAMBERTIAN) {
        const uint32_t mask = __ballot_sync(0xffffffffu, sh.emit);
        if (mask) {
            const int leader = __ffs(mask) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(w.counters + RTW_C_SHADOW, (uint32_t)__popc(mask));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (sh.emit) {
                const uint32_t pos = base + __popc(mask & ((1u << lane) - 1u));
                w.sh_o[pos] = make_float4(sh.o.x, sh.o.y, sh.o.z, sh.tmax);
                w.sh_d[pos] = make_float4(sh.d.x, sh.d.y, sh.d.z, __uint_as_float(slot));
                w.sh_c[pos] = make_float4(sh.contrib.x, sh.contrib.y, sh.contrib.z, __uint_as_float(sh.seed));
            }
        }
    }
}

// All five shading queues in ONE launch.  The queues are cut into 256-entry chunks, numbered
// class by class (lambertian first: the long queue starts first, the short ones fill the tail);
// CTAs take chunks round-robin, so a CTA — and every warp in it — runs one BSDF at a time.  The
// short queues (metal, dielectric, terminal: 5-30 % of the vertices) no longer pay a launch each
// with the machine mostly idle.
__global__ void __launch_bounds__(kBlock, 2) k_shade_all(const DScene sc, const DWave w, uint32_t class_mask) {
    uint32_t n[RTW_Q_COUNT], first_chunk[RTW_Q_COUNT + 1];
    first_chunk[0] = 0;
#pragma unroll
    for (int c = 0; c < RTW_Q_COUNT; c++) {
        n[c] = (class_mask >> c) & 1u ? w.counters[RTW_C_QUEUE0 + c] : 0u;
        first_chunk[c + 1] = first_chunk[c] + (n[c] + kBlock - 1) / kBlock;
    }
    for (uint32_t v = blockIdx.x; v < first_chunk[RTW_Q_COUNT]; v += gridDim.x) {
        if (v < first_chunk[1]) shade_chunk<RTW_Q_LAMBERTIAN>(sc, w, (v - first_chunk[0]) * kBlock, n[0]);
        else if (v < first_chunk[2]) shade_chunk<RTW_Q_METAL>(sc, w, (v - first_chunk[1]) * kBlock, n[1]);
        else if (v < first_chunk[3]) shade_chunk<RTW_Q_DIELECTRIC>(sc, w, (v - first_chunk[2]) * kBlock, n[2]);
        else if (v < first_chunk[4]) shade_chunk<RTW_Q_ISOTROPIC>(sc, w, (v - first_chunk[3]) * kBlock, n[3]);
        else shade_chunk<RTW_Q_TERMINAL>(sc, w, (v - first_chunk[4]) * kBlock, n[4]);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        // ray statistics of this iteration: every queued slot was one closest-hit ray (the shadow
        // rays are counted by k_shadow, once the shadow queue is complete)
        unsigned long long rays = 0;
        for (int c = 0; c < RTW_Q_COUNT; c++) rays += w.counters[RTW_C_QUEUE0 + c];
        atomic_add_u64(w.counters + RTW_C_RAYS_LO, rays);
        w.counters[RTW_C_ALIVE] = (uint32_t)rays;   // 0 ⇒ nothing left in flight: the host stops iterating
    }
}

// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) k_shadow(const DScene sc, const DWave w) {
    extern __shared__ float4 s_top[];
    stage_top(sc, s_top);
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t n = w.counters[RTW_C_SHADOW];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) shadow_thread<MemGlobal>(sc, s_top, w, i);
    if (blockIdx.x == 0 && threadIdx.x == 0) atomic_add_u64(w.counters + RTW_C_SHRAYS_LO, (unsigned long long)n);
}

// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) k_wave_init(const DWave w) {
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x; slot < (uint32_t)w.P; slot += stride) {
        w.misc[slot] = make_uint2(misc_pack(0, RTW_ST_EMPTY), 0u);
        w.queue[(size_t)RTW_Q_REGEN * w.P + slot] = slot;   // every slot is free: the first k_regen fills the pool
    }
    if (blockIdx.x == 0 && threadIdx.x < RTW_C_WORDS)
        w.counters[threadIdx.x] = (threadIdx.x == RTW_C_QUEUE0 + RTW_Q_REGEN) ? (uint32_t)w.P : 0u;
}

// Depth limit 0: `while (depth < maxRayDepth)` (raygen.cu:36) never runs, every sample is
// black; only the sample count reaches the accumulator.
__global__ void __launch_bounds__(kBlock) k_add_count(int n, float count, float4* __restrict__ accum) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) accum[i].w += count;
}

// raygen/raygen.cu:144-158: sqrt gamma of the mean, alpha 1 (Q29)
__global__ void __launch_bounds__(kBlock) k_tonemap(int n, float inv_spp, const float4* __restrict__ accum, float4* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = accum[i];
    out[i] = make_float4(sqrtf(a.x * inv_spp), sqrtf(a.y * inv_spp), sqrtf(a.z * inv_spp), 1.f);
}

// ---- test hooks --------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) k_trace(const DScene sc, const rtw_ray* __restrict__ rays, int n,
                                                   rtw_hit* __restrict__ hits, int32_t* __restrict__ occluded) {
    extern __shared__ float4 s_top[];
    stage_top(sc, s_top);
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const rtw_ray a = rays[i];
        RayW r;
        r.o = f3(a.origin[0], a.origin[1], a.origin[2]);
        r.d = f3(a.dir[0], a.dir[1], a.dir[2]);
        r.tmin = a.tmin; r.tmax = a.tmax; r.time = a.time; r.gather = a.gather_time; r.seed = a.seed;
        if (hits) {
            float t; uint32_t leaf;
            const bool found = traverse_closest<MemGlobal, false>(sc, s_top, r, t, leaf, nullptr);
            rtw_hit h;
            h.t = found ? t : 0.f;
            h.prim = found ? sc.leaf_prim[leaf] : -1;
            hits[i] = h;
        }
        if (occluded) occluded[i] = traverse_any<MemGlobal, false>(sc, s_top, r, nullptr) ? 1 : 0;
    }
}

__global__ void k_rng_stream(uint32_t seed0, uint32_t seed1, int n, uint32_t* tea_out, uint32_t* lcg_out,
                             float* rnd_out, uint32_t* xs_out, float* randf_out) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    const uint32_t s = tea<64>(seed0, seed1);
    if (tea_out) *tea_out = s;
    uint32_t a = s, b = s, c = s, d = s;
    for (int i = 0; i < n; i++) {
        if (lcg_out) lcg_out[i] = lcg(a);
        if (rnd_out) rnd_out[i] = rnd(b);
        if (xs_out) xs_out[i] = xorshift32(c);
        if (randf_out) randf_out[i] = randf(d);
    }
}

__global__ void __launch_bounds__(kBlock) k_math_eval(int fn, const float* __restrict__ x, const float* __restrict__ y, int n,
                                                       float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float a = x[i];
    float r;
    switch (fn) {
        case RTW_FN_SIN: r = det_sinf(a); break;
        case RTW_FN_COS: r = det_cosf(a); break;
        case RTW_FN_LOG: r = det_logf(a); break;
        case RTW_FN_POW5: r = det_pow5f(a); break;
        case RTW_FN_ATAN2: r = det_atan2f(a, y[i]); break;
        default: r = det_asinf(a); break;
    }
    out[i] = r;
}

__global__ void __launch_bounds__(kBlock) k_camera_rays(const DScene sc, const DWave w, rtw_ray* __restrict__ out) {
    const uint32_t npix = (uint32_t)w.Nx * (uint32_t)w.Ny;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    generate_path(sc, w, i, (unsigned long long)i);
    const float4 ro = w.ray_o[i], rd = w.ray_d[i], th = w.thr[i];
    rtw_ray r;
    r.origin[0] = ro.x; r.origin[1] = ro.y; r.origin[2] = ro.z;
    r.dir[0] = rd.x; r.dir[1] = rd.y; r.dir[2] = rd.z;
    r.tmin = 1e-6f; r.tmax = 1.e27f;
    r.time = ro.w; r.gather_time = rd.w;
    r.seed = __float_as_uint(th.w);
    r.pad = 0;
    out[i] = r;
}

template <class K>
int resident_grid(K kernel, size_t smem, int sms) {
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kBlock, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
    return sms * per_sm;
}

}  // namespace

cudaError_t wavefront_configure(const DScene& sc, LaunchCfg* cfg) {
    int dev = 0, sms = 148;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cfg->block = kBlock;
    cfg->smem_top = sizeof(float4) * 4 * (size_t)(sc.n_top > 0 ? sc.n_top : 1);
    if (cfg->smem_top > 48 * 1024) {
        e = cudaFuncSetAttribute(k_traverse, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg->smem_top);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(k_shadow, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg->smem_top);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(k_trace, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg->smem_top);
        if (e != cudaSuccess) return e;
    }
    cfg->grid_logic = resident_grid(k_regen, 0, sms);
    cfg->grid_traverse = resident_grid(k_traverse, cfg->smem_top, sms);
    cfg->grid_shadow = resident_grid(k_shadow, cfg->smem_top, sms);
    cfg->grid_shade = resident_grid(k_shade_all, 0, sms);
    return cudaSuccess;
}

cudaError_t wavefront_init(cudaStream_t s, const DWave& w, uint64_t* launches) {
    k_wave_init<<<296, kBlock, 0, s>>>(w);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

cudaError_t wavefront_iteration(cudaStream_t s, const DScene& sc, const DWave& w, const LaunchCfg& cfg,
                                uint64_t* launches, cudaEvent_t* ev) {
    uint64_t l = 0;
    if (ev) cudaEventRecord(ev[0], s);
    k_regen<<<cfg.grid_logic, kBlock, 0, s>>>(sc, w);
    l++;
    k_advance<<<1, 32, 0, s>>>(w);
    l++;
    if (ev) cudaEventRecord(ev[1], s);
    k_traverse<<<cfg.grid_traverse, kBlock, cfg.smem_top, s>>>(sc, w);
    l++;
    if (ev) cudaEventRecord(ev[2], s);
    k_shade_all<<<cfg.grid_shade, kBlock, 0, s>>>(sc, w, cfg.class_mask | (1u << RTW_Q_TERMINAL));
    l++;
    if (ev) cudaEventRecord(ev[3], s);
    if (cfg.has_shadow) { k_shadow<<<cfg.grid_shadow, kBlock, cfg.smem_top, s>>>(sc, w); l++; }
    if (ev) cudaEventRecord(ev[4], s);
    if (launches) *launches += l;
    return cudaGetLastError();
}

cudaError_t launch_add_count(cudaStream_t s, int n_pixels, int count, float4* accum, uint64_t* launches) {
    k_add_count<<<(n_pixels + kBlock - 1) / kBlock, kBlock, 0, s>>>(n_pixels, (float)count, accum);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

cudaError_t launch_tonemap(cudaStream_t s, int n_pixels, int spp, const float4* accum, float4* out, uint64_t* launches) {
    const float inv = 1.0f / (float)spp;
    k_tonemap<<<(n_pixels + kBlock - 1) / kBlock, kBlock, 0, s>>>(n_pixels, inv, accum, out);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

cudaError_t launch_trace(cudaStream_t s, const DScene& sc, const LaunchCfg& cfg, const rtw_ray* rays, int n,
                         rtw_hit* hits, int32_t* occluded, uint64_t* launches) {
    if (n <= 0) return cudaSuccess;
    int grid = (n + kBlock - 1) / kBlock;
    if (grid > cfg.grid_traverse) grid = cfg.grid_traverse;
    k_trace<<<grid, kBlock, cfg.smem_top, s>>>(sc, rays, n, hits, occluded);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

cudaError_t launch_rng_stream(cudaStream_t s, uint32_t seed0, uint32_t seed1, int n, uint32_t* tea_out,
                              uint32_t* lcg_out, float* rnd_out, uint32_t* xs_out, float* randf_out) {
    k_rng_stream<<<1, 32, 0, s>>>(seed0, seed1, n, tea_out, lcg_out, rnd_out, xs_out, randf_out);
    return cudaGetLastError();
}

cudaError_t launch_math_eval(cudaStream_t s, int fn, const float* x, const float* y, int n, float* out) {
    if (n <= 0) return cudaSuccess;
    k_math_eval<<<(n + kBlock - 1) / kBlock, kBlock, 0, s>>>(fn, x, y, n, out);
    return cudaGetLastError();
}

cudaError_t launch_camera_rays(cudaStream_t s, const DScene& sc, const DWave& w, int sample, rtw_ray* rays_out) {
    DWave t = w;
    t.sample_first = sample;
    t.sample_stride = 0;
    const int npix = w.Nx * w.Ny;
    k_camera_rays<<<(npix + kBlock - 1) / kBlock, kBlock, 0, s>>>(sc, t, rays_out);
    return cudaGetLastError();
}

}  // namespace rtw
