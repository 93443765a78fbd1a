// wavefront.cu — the wavefront path-tracing kernels (sm_100a).
//
// One iteration of the pool (replaces one trip of the `rayColor` while-loop of
// raygen/raygen.cu:36-84 for every live path at once):
//   k_regen       flush the paths that finished last iteration (compact queue), start new camera paths
//                 (its last CTA retires the sample ordinals just handed out and clears the per-iteration counters)
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

// After k_regen's last CTA (before k_traverse): the sample ordinals handed out by k_regen are retired and the
// per-iteration counters (queue lengths, shadow count, rays in flight) are cleared.  One thread.
__device__ __forceinline__ void advance_iteration(const DWave& w) {
    unsigned long long* next = reinterpret_cast<unsigned long long*>(w.counters + RTW_C_NEXT_LO);
    const unsigned long long n = (unsigned long long)w.counters[RTW_C_QUEUE0 + RTW_Q_REGEN];
    if (w.trace_fill) {   // paths k_regen just started (ordinals next .. total - 1 of the n it looked at)
        const unsigned long long left = *next < w.total ? w.total - *next : 0ull;
        w.counters[RTW_C_TRACE0 + w.parity] += (uint32_t)(n < left ? n : left);
        w.counters[RTW_C_TRACE0 + (w.parity ^ 1)] = 0u;   // filled by this iteration's shading for the next one
    }
    *next += n;
    for (int k = 0; k < 8; k++) w.counters[k] = 0u;
    w.counters[RTW_C_WORK_CLOSEST] = 0u;   // work counters of the persistent traversal warps
    w.counters[RTW_C_WORK_SHADOW] = 0u;
    w.counters[RTW_C_TICKET] = 0u;
}

// Regeneration: the slots whose path finished in the previous iteration (a compact
// queue, so every lane of a warp does the same work) are flushed into the accumulator
// and restarted with the next sample ordinal (raygen.cu:129-135, :141-158).
__global__ void __launch_bounds__(kBlock) k_regen(const DScene sc, const DWave w) {
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t n = w.counters[RTW_C_QUEUE0 + RTW_Q_REGEN];
    const unsigned long long next = *reinterpret_cast<const unsigned long long*>(w.counters + RTW_C_NEXT_LO);
    const uint32_t* q = w.queue + (size_t)RTW_Q_REGEN * w.P;
    // sparse mode: the restarted slots follow the survivors that the shading kernel listed (advance_iteration adds the count)
    const uint32_t trace_base = w.trace_fill ? w.counters[RTW_C_TRACE0 + w.parity] : 0u;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        // first iteration of a frame: the pool is empty, slot i is free and there is nothing to flush (no pass
        // over the pool to say so: k_wave_init only sets the counters)
        const uint32_t slot = w.first_iter ? i : q[i];
        // the old path's state is requested first, the new path (tea<64>: ~1000 integer instructions,
        // no memory) is computed while those loads are in flight, then flush, then store
        const float4 ra = w.first_iter ? make_float4(0.f, 0.f, 0.f, u2f(RTW_PIXEL_NONE)) : w.rad[slot];
        // queue position i takes sample ordinal next + i: no atomic (advance_iteration moves `next` on by n
        // after this kernel)
        const unsigned long long ordinal = next + (unsigned long long)i;
        if (ordinal < w.total) {
            const NewPath np = make_path(sc, w, ordinal);
            logic_flush(w, ra);
            store_path(w, slot, np);
            if (w.trace_fill) w.queue[(size_t)RTW_Q_TRACE * w.P + trace_base + i] = slot;
        } else {
            logic_flush(w, ra);
            retire_slot(w, slot);
        }
    }
    // The CTA that finishes last retires the iteration (every other CTA has read the queue length and `next`
    // before it took its ticket): no separate one-warp launch between regeneration and traversal.
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(w.counters + RTW_C_TICKET, 1u) == gridDim.x - 1u) {
            __threadfence();
            advance_iteration(w);
        }
    }
}

// ---------------------------------------------------------------------------
// Persistent traversal warps with dynamic ray fetching.
//
// A warp keeps 32 walkers (device/traverse.h).  All lanes advance by one bounded step
// (<= kStepNodes node visits, then <= kStepPrims primitive tests), so lanes never run
// further apart than one step; a lane whose ray is finished parks its result, and as soon
// as fewer than kMinActive lanes are still walking the warp (converged at that point)
// publishes the parked results with one aggregated push and refills the idle lanes from a
// global work counter.  Incoherent rays (after the first bounce every ray of a warp is in
// a different part of the hierarchy) otherwise leave most lanes waiting for the longest
// walk of the warp: ncu measured 8.3 of 32 lanes active on the Book-2 scene before this.
struct WalkCfg {
    int step_nodes, step_prims, min_active;
    int node_min, prim_min;   // > 0: inner loops under a warp vote (Walker::step_voted)
    // min_active <= 0: static mode — every ray does the same walk (the scene is one leaf range), so
    // warp k simply takes items 32k.., 32(k + #warps).. and no work counter is touched
};

// Job = where rays come from and where results go.
struct ClosestJob {      // stage 1: every live slot of the pool (raygen.cu:41-54)
    static constexpr bool kAny = false;
    const DScene& sc;
    const DWave& w;
    // option sort_rays: only the slots that hold a ray are in `order`
    __device__ uint32_t count() const { return w.use_order ? w.bins[w.n_bins] : (uint32_t)w.P; }
    __device__ uint32_t* work() const { return w.counters + RTW_C_WORK_CLOSEST; }
    // work index -> slot: pool order, or bin by bin
    __device__ uint32_t item_of(uint32_t i) const { return w.use_order ? __ldg(w.order + i) : i; }
    __device__ bool load(uint32_t slot, RayW& r) const {
        const RayRec ray = w.rec[slot].ray;
        const float4 ro = ray.o, rd = ray.d;
        if (f2u(ro.w) == RTW_RAY_NONE) return false;
        r.o = f3(ro);
        r.d = f3(rd);
        r.tmin = 1e-6f;
        r.tmax = 1.e27f;   // shaders/appconfig.h:3
        r.time = ro.w;
        r.gather = rd.w;
        r.seed = sc.has_media ? f2u(w.rec[slot].st.thr.w) : 0u;
        return true;
    }
    // returns the shading class of the finished slot
    __device__ int store(uint32_t slot, bool found, float t, uint32_t leaf) const {
        w.rec[slot].st.hit = make_uint2(f2u(t), found ? leaf : RTW_REF_NONE);
        if (!found) return RTW_Q_TERMINAL;
        const int mt = meta_mattype(f2u(__ldg(sc.prims + 2 * leaf + 1).w));
        if (mt == RTW_MAT_LAMBERTIAN) return RTW_Q_LAMBERTIAN;
        if (mt == RTW_MAT_METAL) return RTW_Q_METAL;
        if (mt == RTW_MAT_DIELECTRIC) return RTW_Q_DIELECTRIC;
        if (mt == RTW_MAT_ISOTROPIC) return RTW_Q_ISOTROPIC;
        return RTW_Q_TERMINAL;
    }
};

struct ShadowJob {       // stage 3: the NEE rays (closehit.cu:16-42,97-113): tmin 500e-7, ray time 0
    static constexpr bool kAny = true;
    const DScene& sc;
    const DWave& w;
    __device__ uint32_t count() const { return w.counters[RTW_C_SHADOW]; }
    __device__ uint32_t* work() const { return w.counters + RTW_C_WORK_SHADOW; }
    __device__ uint32_t item_of(uint32_t i) const { return w.use_order ? __ldg(w.order + i) : i; }
    __device__ bool load(uint32_t i, RayW& r) const {
        const float4 so = w.sh_o[i];
        const float4 sd = w.sh_d[i];
        r.o = f3(so);
        r.d = f3(sd);
        r.tmin = 500 * 1.0e-7f;
        r.tmax = so.w;
        r.time = 0.0f;
        r.gather = sc.has_motion ? w.rec[f2u(sd.w)].ray.d.w : 0.f;
        r.seed = f2u(w.sh_c[i].w);
        return true;
    }
    __device__ int store(uint32_t i, bool occluded, float, uint32_t) const {
        if (!occluded) {
            const float4 c = w.sh_c[i];
            const uint32_t slot = f2u(w.sh_d[i].w);
            float4 ra = w.rad[slot];
            ra.x += c.x; ra.y += c.y; ra.z += c.z;
            w.rad[slot] = ra;
        }
        return -1;
    }
};

struct TraceJob {        // test hook: caller-supplied rays (rtw_trace_closest / rtw_trace_any)
    const DScene& sc;
    const rtw_ray* rays;
    int n;
    rtw_hit* hits;
    int32_t* occluded;
    uint32_t* counter;
    __device__ uint32_t count() const { return (uint32_t)n; }
    __device__ uint32_t* work() const { return counter; }
    __device__ uint32_t item_of(uint32_t i) const { return i; }
    __device__ bool load(uint32_t i, RayW& r) const {
        const rtw_ray a = rays[i];
        r.o = f3(a.origin[0], a.origin[1], a.origin[2]);
        r.d = f3(a.dir[0], a.dir[1], a.dir[2]);
        r.tmin = a.tmin; r.tmax = a.tmax; r.time = a.time; r.gather = a.gather_time; r.seed = a.seed;
        return true;
    }
};
struct TraceClosestJob : TraceJob {
    static constexpr bool kAny = false;
    __device__ int store(uint32_t i, bool found, float t, uint32_t leaf) const {
        rtw_hit h;
        h.t = found ? t : 0.f;
        h.prim = found ? sc.leaf_prim[leaf] : -1;
        hits[i] = h;
        return -1;
    }
};
struct TraceAnyJob : TraceJob {
    static constexpr bool kAny = true;
    __device__ int store(uint32_t i, bool found, float, uint32_t) const {
        occluded[i] = found ? 1 : 0;
        return -1;
    }
};

template <class Job, bool kPush>
__device__ __forceinline__ void walk_persistent(const DScene& sc, const DWave* wq, const float4* s_top, const Job& job,
                                                const WalkCfg wc) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lt = (1u << lane) - 1u;
    const uint32_t total = job.count();
    Walker<MemGlobal, Job::kAny> wk;
    uint32_t stack[RTW_STACK];
    wk.done = true;
    bool active = false;
    uint32_t item = 0;
    int parked_cls = -1;         // shading class of a finished slot waiting for the aggregated push
    uint32_t parked_slot = 0;
    const bool static_work = wc.min_active <= 0;
    uint32_t static_base = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 32u;
    const uint32_t static_stride = gridDim.x * blockDim.x;
    for (;;) {
        // ---- converged: publish parked results, refill idle lanes ----
        if (kPush) {
            const uint32_t peers = __match_any_sync(0xffffffffu, parked_cls);
            const int leader = __ffs(peers) - 1;
            uint32_t base = 0;
            if (parked_cls >= 0 && (int)lane == leader)
                base = atomicAdd(wq->counters + RTW_C_QUEUE0 + parked_cls, (uint32_t)__popc(peers));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (parked_cls >= 0) wq->queue[(size_t)parked_cls * wq->P + base + __popc(peers & lt)] = parked_slot;
            parked_cls = -1;
        }
        const uint32_t idle = __ballot_sync(0xffffffffu, !active);
        uint32_t base = 0;
        if (static_work) {
            base = static_base;
            static_base += static_stride;
        } else {
            if (lane == 0) base = atomicAdd(job.work(), (uint32_t)__popc(idle));
            base = __shfl_sync(0xffffffffu, base, 0);
        }
        if (!active) {
            const uint32_t idx = base + __popc(idle & lt);
            if (idx < total) {
                RayW r;
                const uint32_t it = job.item_of(idx);   // work index -> item (ClosestJob: the slot, in pool or sorted order)
                if (job.load(it, r)) {
                    wk.begin(sc, r, stack);
                    active = true;
                    item = it;
                }
            }
        }
        if (!__any_sync(0xffffffffu, active)) {
            if (base >= total) break;   // nothing walking, nothing left to fetch
            continue;                   // the fetched items needed no traversal (empty slots): fetch again
        }
        // ---- walk until too few lanes are still busy (or, once the work is gone, until all are done) ----
        const int min_active = (!static_work && base + 32u <= total) ? wc.min_active : 1;
        do {
            if (wc.node_min > 0) wk.step_voted(sc, s_top, stack, wc.step_nodes, wc.step_prims, wc.node_min, wc.prim_min);
            else if (active) wk.step(sc, s_top, stack, wc.step_nodes, wc.step_prims);
            if (active) {
                if (wk.done) {
                    const int cls = job.store(item, wk.found, wk.tb, wk.lb);
                    if (kPush) { parked_cls = cls; parked_slot = item; }
                    active = false;
                }
            }
            __syncwarp();
        } while (__popc(__ballot_sync(0xffffffffu, active)) >= min_active);
    }
}

// Flat variant for scenes that are ONE leaf range (<= leaf_size primitives, e.g. the Cornell box):
// every ray walks the same primitive list, so lanes stay in lockstep by construction and the
// straight-line per-thread traversal (device/traverse.h) beats the resumable walker by ~20 %.
__global__ void __launch_bounds__(kBlock) k_traverse_flat(const DScene sc, const DWave w) {
    extern __shared__ float4 s_top[];
    if (!(sc.root & RTW_REF_LEAF) && !sc.wide) stage_top(sc, s_top);
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x; slot - lane < (uint32_t)w.P; slot += stride) {
        int cls = -1;
        if (slot < (uint32_t)w.P) cls = traverse_thread<MemGlobal>(sc, s_top, w, slot);
        __syncwarp();
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

__global__ void __launch_bounds__(kBlock) k_shadow_flat(const DScene sc, const DWave w) {
    extern __shared__ float4 s_top[];
    if (!(sc.root & RTW_REF_LEAF) && !sc.wide) stage_top(sc, s_top);
    const uint32_t stride = gridDim.x * blockDim.x;
    const uint32_t n = w.counters[RTW_C_SHADOW];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) shadow_thread<MemGlobal>(sc, s_top, w, i);
    if (blockIdx.x == 0 && threadIdx.x == 0) atomic_add_u64(w.counters + RTW_C_SHRAYS_LO, (unsigned long long)n);
}

// Small-scene variants (layout.h SmallScene): the leaf records arrive as a kernel parameter, the
// loop over them is warp-uniform.  Same per-slot work as the flat kernels otherwise.
// Control flow is kept warp-convergent around the walk (whole 256-slot tiles, a slot without a
// live ray walks a null ray and drops the result), which is what lets the compiler keep the leaf
// loop on the uniform datapath.
__global__ void __launch_bounds__(kBlock) k_traverse_small(const DScene sc, const DWave w, const __grid_constant__ SmallScene ss) {
    const uint32_t lane = threadIdx.x & 31u;
    // the whole pool in slot order, or (sparse mode) the list of slots that hold a ray
    const uint32_t n = w.trace_use ? w.counters[RTW_C_TRACE0 + w.parity] : (uint32_t)w.P;
    const uint32_t* list = w.queue + (size_t)RTW_Q_TRACE * w.P;
    const uint32_t tiles = (n + kBlock - 1) / kBlock;
    for (uint32_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        uint32_t slot = tile * kBlock + threadIdx.x;
        RayW r;
        r.o = f3(0.f); r.d = f3(0.f); r.tmin = 1e-6f; r.tmax = 1.e27f; r.time = 0.f; r.gather = 0.f; r.seed = 0u;
        bool alive = false;
        if (slot < n) {
            if (w.trace_use) slot = list[slot];
            alive = closest_ray_of(sc, w, slot, r);
        }
        // a slot without a ray (drain tail of a frame) walks a harmless ray: the zero ray of a retired slot
        // would send every division of its warp (k / 0) through the out-of-line slow path
        if (!alive) { r.o = f3(0.f); r.d = f3(1.f); }
        // warp-uniform: a warp of empty slots walks nothing.  (Packing the live rays of a tile into fewer warps
        // was tried for the drain tail of a frame and is slower: one warp's walk is a ~1400-instruction
        // dependent chain, so a tile takes as long with one walking warp as with eight.)
        if (!__any_sync(0xffffffffu, alive)) continue;
        float t;
        uint32_t leaf;
        const bool found = small_closest(ss, r, t, leaf);
        int cls = -1;
        if (alive) cls = closest_result<MemGlobal>(sc, w, slot, found, t, leaf);
        __syncwarp();
        const uint32_t peers = __match_any_sync(0xffffffffu, cls);
        const int leader = __ffs(peers) - 1;
        uint32_t base = 0;
        if (cls >= 0 && (int)lane == leader) base = atomicAdd(w.counters + RTW_C_QUEUE0 + cls, (uint32_t)__popc(peers));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (cls >= 0) w.queue[(size_t)cls * w.P + base + __popc(peers & ((1u << lane) - 1u))] = slot;
    }
}

__global__ void __launch_bounds__(kBlock) k_shadow_small(const DScene sc, const DWave w, const __grid_constant__ SmallScene ss) {
    const uint32_t n = w.counters[RTW_C_SHADOW];
    const uint32_t tiles = (n + kBlock - 1) / kBlock;
    uint32_t n_unocc = 0;
    for (uint32_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const uint32_t i = tile * kBlock + threadIdx.x;
        RayW r;
        r.o = f3(0.f); r.d = f3(1.f); r.tmin = 500 * 1.0e-7f; r.tmax = 0.f; r.time = 0.f; r.gather = 0.f; r.seed = 0u;
        float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
        uint32_t slot = 0;
        const bool valid = i < n;
        if (valid) shadow_ray_of(sc, w, i, r, c, slot);
        const bool occluded = small_any(ss, r) || !valid;
        if (valid) shadow_result(w, slot, c, occluded);
        n_unocc += (uint32_t)__popc(__ballot_sync(0xffffffffu, valid && !occluded));
    }
    if ((threadIdx.x & 31u) == 0u && n_unocc) atomic_add_u64(w.counters + RTW_C_UNOCC_LO, (unsigned long long)n_unocc);
    if (blockIdx.x == 0 && threadIdx.x == 0) atomic_add_u64(w.counters + RTW_C_SHRAYS_LO, (unsigned long long)n);
}

// ---- option sort_rays: counting sort of a stage's rays by (origin cell, direction bin) -----------------------------
// Incoherent rays leave the lanes of a walker warp in different parts of the hierarchy (ncu, scene 4: 8.8 of 32 lanes
// per instruction).  Rays that start in the same cell and head the same way walk the same nodes: the stage takes its
// rays bin by bin.  Which ray a lane gets does not change the ray's result (closest hit = min (t, primitive id)).
//   k_ray_bin    item -> key; rank inside the bin from ONE atomicAdd per warp and key (__match_any_sync)
//   k_bin_sum, k_bin_scan   bins: counts -> first positions; bins[n_bins] = items placed
//   k_ray_place  order[first[key] + rank] = item
__device__ __forceinline__ uint32_t ray_key(const DWave& w, float3 o, float3 d) {
    const float top = (float)((1 << w.key_obits) - 1);
    const uint32_t ix = (uint32_t)fminf(fmaxf((o.x - w.key_lo.x) * w.key_scale.x, 0.f), top);
    const uint32_t iy = (uint32_t)fminf(fmaxf((o.y - w.key_lo.y) * w.key_scale.y, 0.f), top);
    const uint32_t iz = (uint32_t)fminf(fmaxf((o.z - w.key_lo.z) * w.key_scale.z, 0.f), top);
    uint32_t m = 0;
    for (int b = 0; b < w.key_obits; b++) m |= (((ix >> b) & 1u) << (3 * b + 2)) | (((iy >> b) & 1u) << (3 * b + 1)) | (((iz >> b) & 1u) << (3 * b));
    const float dm = fmaxf(fmaxf(fabsf(d.x), fabsf(d.y)), fmaxf(fabsf(d.z), 1e-30f));
    const float half = (float)(1 << (w.key_dbits - 1)), dtop = (float)((1 << w.key_dbits) - 1);
    const float rdm = __frcp_rn(dm);   // (a key is a heuristic: no IEEE divisions here)
    const uint32_t qx = (uint32_t)fminf(fmaxf((d.x * rdm + 1.0f) * half, 0.f), dtop);
    const uint32_t qy = (uint32_t)fminf(fmaxf((d.y * rdm + 1.0f) * half, 0.f), dtop);
    const uint32_t qz = (uint32_t)fminf(fmaxf((d.z * rdm + 1.0f) * half, 0.f), dtop);
    return (m << (3 * w.key_dbits)) | (qx << (2 * w.key_dbits)) | (qy << w.key_dbits) | qz;
}
// Camera rays (depth 0) all leave one point: their cell says nothing and an octant holds the whole image, so they get
// bins of their own behind the regular ones, by DIRECTION on a cube map — face, then 256 x 256 cells per face.
// Neighbouring pixels then sit in the same or the next bin whatever slots their paths were started in.
__device__ __forceinline__ uint32_t camera_key(float3 d) {
    const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    uint32_t face;
    float m, u, v;
    if (ax >= ay && ax >= az) { face = d.x < 0.f ? 1u : 0u; m = ax; u = d.y; v = d.z; }
    else if (ay >= az) { face = d.y < 0.f ? 3u : 2u; m = ay; u = d.x; v = d.z; }
    else { face = d.z < 0.f ? 5u : 4u; m = az; u = d.x; v = d.y; }
    const float rm = __frcp_rn(fmaxf(m, 1e-30f));
    const uint32_t iu = (uint32_t)fminf(fmaxf((u * rm + 1.0f) * 128.0f, 0.f), 255.f);
    const uint32_t iv = (uint32_t)fminf(fmaxf((v * rm + 1.0f) * 128.0f, 0.f), 255.f);
    return (face << 16) | (iu << 8) | iv;
}
// Four items per thread (i, i + 256, i + 512, i + 768 of the CTA's 1024): the kernel is bound by memory latency
// (ncu: long-scoreboard 41 warps per issue with one item per thread), so all loads are issued before any key is formed.
constexpr int kBinItems = 4;
template <bool kShadow>
__global__ void __launch_bounds__(kBlock) k_ray_bin(const DWave w) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t n = kShadow ? w.counters[RTW_C_SHADOW] : (uint32_t)w.P;
    const uint32_t first = blockIdx.x * (kBlock * kBinItems) + threadIdx.x;
    if (first - lane >= n) return;   // the CTA's first row of warps is past the end: so are the others
    float4 o[kBinItems], d[kBinItems];
    uint32_t depth[kBinItems];
#pragma unroll
    for (int k = 0; k < kBinItems; k++) {
        const uint32_t i = first + k * kBlock;
        o[k] = make_float4(0.f, 0.f, 0.f, u2f(RTW_RAY_NONE));
        d[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        depth[k] = 1u;
        if (i < n) {
            if (kShadow) {
                o[k] = w.sh_o[i];
                d[k] = w.sh_d[i];
            } else {
                const RayRec ray = w.rec[i].ray;
                o[k] = ray.o;
                d[k] = ray.d;
                depth[k] = misc_depth(w.rec[i].st.misc.x);   // (the same 64-byte line)
            }
        }
    }
#pragma unroll
    for (int k = 0; k < kBinItems; k++) {
        const uint32_t i = first + k * kBlock;
        if (i - lane >= n) break;        // whole warps only: the vote below needs every lane
        uint32_t key = 0x80000000u | lane;   // no ray: a key of its own (matches no other lane), not binned
        if (i < n && (kShadow || f2u(o[k].w) != RTW_RAY_NONE)) {
            const bool primary = !kShadow && w.key_camera && depth[k] == 0u;
            key = primary ? (1u << (3 * (w.key_obits + w.key_dbits))) + camera_key(f3(d[k])) : ray_key(w, f3(o[k]), f3(d[k]));
        }
        const uint32_t peers = __match_any_sync(0xffffffffu, key);
        const int leader = __ffs(peers) - 1;
        uint32_t base = 0;
        const bool binned = !(key & 0x80000000u);
        if (binned && (int)lane == leader) base = atomicAdd(w.bins + key, (uint32_t)__popc(peers));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (i < n) w.bin_rank[i] = make_uint2(key, base + (uint32_t)__popc(peers & ((1u << lane) - 1u)));
    }
}
// bins: counts -> first positions, in two launches of n_bins / 1024 CTAs (a one-CTA scan of 2^18 bins is a chain of
// 512 dependent L2 accesses per thread: 0.5 ms).  part[] = the sums of the 1024-bin chunks, behind bins[n_bins].
__device__ __forceinline__ uint32_t block_scan_1024(uint32_t v, uint32_t* s_warp, uint32_t& total) {   // exclusive
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += u;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t before = 0, all = 0;
#pragma unroll
    for (int k = 0; k < 32; k++) {
        const uint32_t x = s_warp[k];
        before += k < warp ? x : 0u;
        all += x;
    }
    total = all;
    return before + inc - v;
}
__global__ void __launch_bounds__(1024) k_bin_sum(const DWave w) {
    __shared__ uint32_t s_warp[32];
    const int n_bins = w.n_bins;
    const int k = blockIdx.x * 1024 + threadIdx.x;
    uint32_t total;
    block_scan_1024(k < n_bins ? w.bins[k] : 0u, s_warp, total);
    if (threadIdx.x == 0) w.bins[n_bins + 1 + blockIdx.x] = total;
}
__global__ void __launch_bounds__(1024) k_bin_scan(const DWave w) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_warp2[32];
    const int n_bins = w.n_bins;
    const uint32_t* part = w.bins + n_bins + 1;
    uint32_t mine = 0;   // chunks in front of this one (at most 256)
    for (int j = threadIdx.x; j < (int)blockIdx.x; j += 1024) mine += part[j];
    uint32_t base;
    block_scan_1024(mine, s_warp2, base);
    const int k = blockIdx.x * 1024 + threadIdx.x;
    const uint32_t c = k < n_bins ? w.bins[k] : 0u;
    uint32_t total;
    const uint32_t excl = block_scan_1024(c, s_warp, total);
    if (k < n_bins) w.bins[k] = base + excl;
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) w.bins[n_bins] = base + total;   // items placed
}
template <bool kShadow>
__global__ void __launch_bounds__(kBlock) k_ray_place(const DWave w) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n = kShadow ? w.counters[RTW_C_SHADOW] : (uint32_t)w.P;
    if (i >= n) return;
    const uint2 kr = w.bin_rank[i];
    if (!(kr.x & 0x80000000u)) w.order[w.bins[kr.x] + kr.y] = i;
}
// One counting sort on stream s (a memset of the bins + 4 launches).
template <bool kShadow>
cudaError_t bin_rays(cudaStream_t s, const DWave& w, uint64_t* launches) {
    const int n_bins = w.n_bins;
    cudaError_t e = cudaMemsetAsync(w.bins, 0, sizeof(uint32_t) * (size_t)(n_bins + 1), s);
    if (e != cudaSuccess) return e;
    const int grid = (w.P + kBlock - 1) / kBlock;
    k_ray_bin<kShadow><<<(w.P + kBlock * kBinItems - 1) / (kBlock * kBinItems), kBlock, 0, s>>>(w);
    k_bin_sum<<<(n_bins + 1023) / 1024, 1024, 0, s>>>(w);
    k_bin_scan<<<(n_bins + 1023) / 1024, 1024, 0, s>>>(w);
    k_ray_place<kShadow><<<grid, kBlock, 0, s>>>(w);
    if (launches) *launches += 4;
    return cudaGetLastError();
}

// resident CTAs per SM the persistent walkers' registers are held to: 4 -> 64 registers, 24-90 bytes of spills.
// Measured (profiles/r02/lbvh_r02.md): scene 4 432 -> 453 Mpaths/s, C5 204 -> 210 against 3 CTAs / 76 registers.
#ifndef RTW_WALK_CTAS
#define RTW_WALK_CTAS 4
#endif
__global__ void __launch_bounds__(kBlock, RTW_WALK_CTAS) k_traverse(const DScene sc, const DWave w, const WalkCfg wc) {
    extern __shared__ float4 s_top[];
    stage_top(sc, s_top);
    const ClosestJob job{sc, w};
    walk_persistent<ClosestJob, true>(sc, &w, s_top, job, wc);
}

// ---------------------------------------------------------------------------
// One 256-wide chunk [i0, i0 + 256) of shading queue kClass: every warp runs ONE BSDF.
// kFuse (small scenes, option fuse_shadow): the NEE shadow ray is traced HERE, against the by-value copy of the
// scene, instead of travelling through the shadow queue to k_shadow_small — 96 B of DRAM traffic per NEE ray
// (48 B written + 48 B read) and one launch per iteration less.  The any-hit walk is warp-uniform (every lane
// walks the same 13 records; a lane without an NEE ray walks a null ray), the unoccluded contribution goes
// into the slot's radiance record exactly as k_shadow_small would have added it, in the same order.
template <int kClass, bool kFuse>
__device__ __forceinline__ void shade_chunk(const DScene& sc, const DWave& w, uint32_t slot, uint32_t ahead, const SmallScene* ss,
                                            uint32_t& n_shadow) {
    const uint32_t lane = threadIdx.x & 31u;
    ShadowOut sh;
    sh.emit = false;
    bool finished = false;
    if (slot != RTW_REF_NONE) finished = shade_thread<kClass, MemGlobal>(sc, w, slot, sh, ahead);
    else prefetch_slot(w, ahead);
    __syncwarp();
    // The warp's pushes — finished slots (regen queue), NEE rays (shadow queue), survivors (sparse-mode
    // trace list) — reserve their space with ONE atomic round trip: lanes 0, 1, 2 each add one count.
    const bool survivor = w.trace_fill && slot != RTW_REF_NONE && !finished;
    const uint32_t m_regen = __ballot_sync(0xffffffffu, finished);
    uint32_t m_shadow = kClass == RTW_Q_LAMBERTIAN ? __ballot_sync(0xffffffffu, sh.emit) : 0u;
    const uint32_t m_trace = w.trace_fill ? __ballot_sync(0xffffffffu, survivor) : 0u;
    if (kFuse && kClass == RTW_Q_LAMBERTIAN && m_shadow != 0u) {
        RayW r;   // traceOcclusion, shaders/closehit.cu:16-42,97-113: tmin 500e-7, ray time 0
        r.o = f3(0.f); r.d = f3(1.f); r.tmin = 500 * 1.0e-7f; r.tmax = 0.f; r.time = 0.f; r.gather = 0.f; r.seed = 0u;
        if (sh.emit) {
            r.o = sh.o; r.d = sh.d; r.tmax = sh.tmax;
            r.gather = sc.has_motion ? sh.gather : 0.f;
            r.seed = sh.seed;
        }
        const bool occluded = small_any(*ss, r);
        const bool add = sh.emit && !occluded;
        if (add) {
            float4 ra = w.rad[slot];
            ra.x += sh.contrib.x; ra.y += sh.contrib.y; ra.z += sh.contrib.z;
            w.rad[slot] = ra;
        }
        // every lane keeps the warp's counts (NEE rays in the low half, contributions added in the high half);
        // lane 0 publishes them once per kernel
        n_shadow += (uint32_t)__popc(m_shadow) + ((uint32_t)__popc(__ballot_sync(0xffffffffu, add)) << 16);
        if ((n_shadow & 0xffffu) >= 0x8000u) {   // keep the two 16-bit halves from running into each other
            if (lane == 0) {
                atomic_add_u64(w.counters + RTW_C_SHRAYS_LO, (unsigned long long)(n_shadow & 0xffffu));
                atomic_add_u64(w.counters + RTW_C_UNOCC_LO, (unsigned long long)(n_shadow >> 16));
            }
            n_shadow = 0u;
        }
        m_shadow = 0u;
    }
    if ((m_regen | m_shadow | m_trace) == 0u) return;
    const uint32_t mine = lane == 0 ? m_regen : (lane == 1 ? m_shadow : m_trace);
    const int counter = lane == 0 ? RTW_C_QUEUE0 + RTW_Q_REGEN : (lane == 1 ? RTW_C_SHADOW : RTW_C_TRACE0 + (w.parity ^ 1));
    uint32_t base = 0;
    if (lane < 3 && mine) base = atomicAdd(w.counters + counter, (uint32_t)__popc(mine));
    const uint32_t b_regen = __shfl_sync(0xffffffffu, base, 0);
    const uint32_t b_shadow = __shfl_sync(0xffffffffu, base, 1);
    const uint32_t b_trace = __shfl_sync(0xffffffffu, base, 2);
    const uint32_t below = (1u << lane) - 1u;
    if (finished) w.queue[(size_t)RTW_Q_REGEN * w.P + b_regen + __popc(m_regen & below)] = slot;
    if (survivor) w.queue[(size_t)RTW_Q_TRACE * w.P + b_trace + __popc(m_trace & below)] = slot;
    if (!kFuse && kClass == RTW_Q_LAMBERTIAN && sh.emit) {
        const uint32_t pos = b_shadow + __popc(m_shadow & below);
        w.sh_o[pos] = make_float4(sh.o.x, sh.o.y, sh.o.z, sh.tmax);
        w.sh_d[pos] = make_float4(sh.d.x, sh.d.y, sh.d.z, __uint_as_float(slot));
        w.sh_c[pos] = make_float4(sh.contrib.x, sh.contrib.y, sh.contrib.z, __uint_as_float(sh.seed));
    }
}

// All five shading queues in ONE launch.  The queues are cut into 256-entry chunks, numbered
// class by class (lambertian first: the long queue starts first, the short ones fill the tail);
// CTAs take chunks round-robin, so a CTA — and every warp in it — runs one BSDF at a time.  The
// short queues (metal, dielectric, terminal: 5-30 % of the vertices) no longer pay a launch each
// with the machine mostly idle.
// kCtas = resident CTAs per SM the register allocation is held to (option `shade_ctas`): 2 → 95 registers,
// 3 → 80, 4 → 64; the kernel is bound by memory latency, so warps in flight are what it is short of.
template <bool kFuse>
__device__ __forceinline__ void shade_all_body(const DScene& sc, const DWave& w, uint32_t class_mask, const SmallScene* ss) {
    // queue lengths and the first chunk number of every class: shared memory, not registers (the kernel is
    // register-bound: 128 with the tables in registers and spills, see profiles/sass/INDEX.md)
    __shared__ uint32_t s_n[RTW_Q_COUNT], s_first[RTW_Q_COUNT + 1];
    if (threadIdx.x == 0) {
        uint32_t f = 0;
        for (int c = 0; c < RTW_Q_COUNT; c++) {
            const uint32_t n = (class_mask >> c) & 1u ? w.counters[RTW_C_QUEUE0 + c] : 0u;
            s_n[c] = n;
            s_first[c] = f;
            f += (n + kBlock - 1) / kBlock;
        }
        s_first[RTW_Q_COUNT] = f;
    }
    __syncthreads();
    const uint32_t chunks = s_first[RTW_Q_COUNT];
    // class of chunk v
    auto class_of = [&](uint32_t v) -> int {
        int c = 0;
#pragma unroll
        for (int k = 1; k < RTW_Q_COUNT; k++) c += (v >= s_first[k]) ? 1 : 0;
        return c;
    };
    // this thread's queue entry of chunk v (RTW_REF_NONE past the end of its queue)
    auto entry = [&](uint32_t v) -> uint32_t {
        if (v >= chunks) return RTW_REF_NONE;
        const int c = class_of(v);
        const uint32_t i = (v - s_first[c]) * kBlock + threadIdx.x;
        return i < s_n[c] ? w.queue[(size_t)c * w.P + i] : RTW_REF_NONE;
    };
    // Two chunks of queue entries are in flight: `ahead` (loaded one trip ago) addresses the record prefetch
    // inside the chunk, `ahead2` is requested now — so the prefetch never waits for its own index load.
    uint32_t slot = entry(blockIdx.x);
    uint32_t ahead = entry(blockIdx.x + gridDim.x);
    uint32_t n_shadow = 0;   // kFuse: NEE rays this warp traced in place
    for (uint32_t v = blockIdx.x; v < chunks; v += gridDim.x) {
        const uint32_t ahead2 = entry(v + 2 * gridDim.x);
        const int c = class_of(v);
        if (c == RTW_Q_LAMBERTIAN) shade_chunk<RTW_Q_LAMBERTIAN, kFuse>(sc, w, slot, ahead, ss, n_shadow);
        else if (c == RTW_Q_METAL) shade_chunk<RTW_Q_METAL, kFuse>(sc, w, slot, ahead, ss, n_shadow);
        else if (c == RTW_Q_DIELECTRIC) shade_chunk<RTW_Q_DIELECTRIC, kFuse>(sc, w, slot, ahead, ss, n_shadow);
        else if (c == RTW_Q_ISOTROPIC) shade_chunk<RTW_Q_ISOTROPIC, kFuse>(sc, w, slot, ahead, ss, n_shadow);
        else shade_chunk<RTW_Q_TERMINAL, kFuse>(sc, w, slot, ahead, ss, n_shadow);
        slot = ahead;
        ahead = ahead2;
    }
    if (kFuse && (threadIdx.x & 31u) == 0u && n_shadow) {
        atomic_add_u64(w.counters + RTW_C_SHRAYS_LO, (unsigned long long)(n_shadow & 0xffffu));
        atomic_add_u64(w.counters + RTW_C_UNOCC_LO, (unsigned long long)(n_shadow >> 16));
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
template <int kCtas>
__global__ void __launch_bounds__(kBlock, kCtas) k_shade_all(const DScene sc, const DWave w, uint32_t class_mask) {
    shade_all_body<false>(sc, w, class_mask, nullptr);
}
// small scenes with NEE: the shadow rays are traced in place (shade_chunk, kFuse)
template <int kCtas>
__global__ void __launch_bounds__(kBlock, kCtas) k_shade_all_fused(const DScene sc, const DWave w, uint32_t class_mask,
                                                                    const __grid_constant__ SmallScene ss) {
    shade_all_body<true>(sc, w, class_mask, &ss);
}

// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock, RTW_WALK_CTAS) k_shadow(const DScene sc, const DWave w, const WalkCfg wc) {
    extern __shared__ float4 s_top[];
    stage_top(sc, s_top);
    const ShadowJob job{sc, w};
    walk_persistent<ShadowJob, false>(sc, nullptr, s_top, job, wc);
    if (blockIdx.x == 0 && threadIdx.x == 0)
        atomic_add_u64(w.counters + RTW_C_SHRAYS_LO, (unsigned long long)w.counters[RTW_C_SHADOW]);
}

// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) k_wave_init(const DWave w) {
    // every slot is free: the first k_regen (DWave::first_iter) fills slots 0 .. P-1 and retires those past the last sample
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
                                                   rtw_hit* __restrict__ hits, int32_t* __restrict__ occluded,
                                                   uint32_t* __restrict__ work /* two zeroed words */, const WalkCfg wc) {
    extern __shared__ float4 s_top[];
    stage_top(sc, s_top);
    if (hits) {
        const TraceClosestJob job{{sc, rays, n, hits, occluded, work}};
        walk_persistent<TraceClosestJob, false>(sc, nullptr, s_top, job, wc);
    }
    if (occluded) {
        const TraceAnyJob job{{sc, rays, n, hits, occluded, work + 1}};
        walk_persistent<TraceAnyJob, false>(sc, nullptr, s_top, job, wc);
    }
}

__global__ void __launch_bounds__(kBlock) k_trace_small(const __grid_constant__ SmallScene ss, const rtw_ray* __restrict__ rays, int n,
                                                         rtw_hit* __restrict__ hits, int32_t* __restrict__ occluded) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = i < n;
    RayW r;
    r.o = f3(0.f); r.d = f3(1.f); r.tmin = 0.f; r.tmax = 0.f; r.time = 0.f; r.gather = 0.f; r.seed = 0u;
    if (valid) {
        const rtw_ray q = rays[i];
        r.o = f3(q.origin[0], q.origin[1], q.origin[2]);
        r.d = f3(q.dir[0], q.dir[1], q.dir[2]);
        r.tmin = q.tmin; r.tmax = q.tmax; r.time = q.time; r.gather = q.gather_time; r.seed = q.seed;
    }
    if (hits) {
        float t;
        uint32_t leaf;
        int32_t prim;
        const bool found = small_closest(ss, r, t, leaf, &prim);
        if (valid) {
            rtw_hit h;
            h.t = found ? t : 0.f;
            h.prim = prim;
            hits[i] = h;
        }
    }
    if (occluded) {
        const bool occ = small_any(ss, r) || !valid;
        if (valid) occluded[i] = occ ? 1 : 0;
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
    const NewPath np = make_path(sc, w, (unsigned long long)i);
    const float4 ro = np.ray.o, rd = np.ray.d, th = np.st.thr;
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
        e = cudaFuncSetAttribute(k_traverse_flat, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg->smem_top);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(k_shadow_flat, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cfg->smem_top);
        if (e != cudaSuccess) return e;
    }
    cfg->grid_logic = resident_grid(k_regen, 0, sms);
    if (cfg->small) {
        cfg->grid_traverse = resident_grid(k_traverse_small, 0, sms);
        cfg->grid_shadow = resident_grid(k_shadow_small, 0, sms);
    } else if (!cfg->walker) {
        cfg->grid_traverse = resident_grid(k_traverse_flat, cfg->smem_top, sms);
        cfg->grid_shadow = resident_grid(k_shadow_flat, cfg->smem_top, sms);
    } else {
        cfg->grid_traverse = resident_grid(k_traverse, cfg->smem_top, sms);
        cfg->grid_shadow = resident_grid(k_shadow, cfg->smem_top, sms);
    }
    if (cfg->fuse_shadow && cfg->small)
        cfg->grid_shade = cfg->shade_ctas == 2 ? resident_grid(k_shade_all_fused<2>, 0, sms)
                        : cfg->shade_ctas == 4 ? resident_grid(k_shade_all_fused<4>, 0, sms) : resident_grid(k_shade_all_fused<3>, 0, sms);
    else
        cfg->grid_shade = cfg->shade_ctas == 2 ? resident_grid(k_shade_all<2>, 0, sms)
                        : cfg->shade_ctas == 4 ? resident_grid(k_shade_all<4>, 0, sms) : resident_grid(k_shade_all<3>, 0, sms);
    return cudaSuccess;
}

cudaError_t wavefront_init(cudaStream_t s, const DWave& w, uint64_t* launches) {
    k_wave_init<<<1, kBlock, 0, s>>>(w);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

cudaError_t wavefront_iteration(cudaStream_t s, const DScene& sc, const DWave& w, const LaunchCfg& cfg,
                                uint64_t* launches, cudaEvent_t* ev) {
    uint64_t l = 0;
    if (ev) cudaEventRecord(ev[0], s);
    k_regen<<<cfg.grid_logic, kBlock, 0, s>>>(sc, w);
    l++;
    if (ev) cudaEventRecord(ev[1], s);
    if (cfg.small) k_traverse_small<<<cfg.grid_traverse, kBlock, 0, s>>>(sc, w, *cfg.small);
    else if (!cfg.walker) k_traverse_flat<<<cfg.grid_traverse, kBlock, cfg.smem_top, s>>>(sc, w);
    else {
        DWave wt = w;
        wt.use_order = 0;
        if (w.bins) {   // option sort_rays: the walkers take the rays bin by bin
            const cudaError_t e = bin_rays<false>(s, w, &l);
            if (e != cudaSuccess) return e;
            wt.use_order = 1;
            if (ev) cudaEventRecord(ev[1], s);   // the binning is booked on the regeneration stage: `traverse` stays the walk itself
        }
        k_traverse<<<cfg.grid_traverse, kBlock, cfg.smem_top, s>>>(sc, wt, WalkCfg{cfg.step_nodes, cfg.step_prims, cfg.min_active, cfg.node_min, cfg.prim_min});
    }
    l++;
    if (cfg.mid_event) cudaEventRecord(cfg.mid_event, s);
    if (ev) cudaEventRecord(ev[2], s);
    const uint32_t classes = cfg.class_mask | (1u << RTW_Q_TERMINAL);
    const bool fused = cfg.fuse_shadow && cfg.small && cfg.has_shadow;
    if (fused) {
        if (cfg.shade_ctas == 2) k_shade_all_fused<2><<<cfg.grid_shade, kBlock, 0, s>>>(sc, w, classes, *cfg.small);
        else if (cfg.shade_ctas == 4) k_shade_all_fused<4><<<cfg.grid_shade, kBlock, 0, s>>>(sc, w, classes, *cfg.small);
        else k_shade_all_fused<3><<<cfg.grid_shade, kBlock, 0, s>>>(sc, w, classes, *cfg.small);
    } else if (cfg.shade_ctas == 2) k_shade_all<2><<<cfg.grid_shade, kBlock, 0, s>>>(sc, w, classes);
    else if (cfg.shade_ctas == 4) k_shade_all<4><<<cfg.grid_shade, kBlock, 0, s>>>(sc, w, classes);
    else k_shade_all<3><<<cfg.grid_shade, kBlock, 0, s>>>(sc, w, classes);
    l++;
    if (ev) cudaEventRecord(ev[3], s);
    if (cfg.has_shadow && !fused) {
        if (cfg.small) k_shadow_small<<<cfg.grid_shadow, kBlock, 0, s>>>(sc, w, *cfg.small);
        else if (!cfg.walker) k_shadow_flat<<<cfg.grid_shadow, kBlock, cfg.smem_top, s>>>(sc, w);
        else {
            DWave wt = w;
            wt.use_order = 0;
            if (w.bins && cfg.sort_shadow) {   // the NEE rays too (the queue is in the order the shading warps finished)
                const cudaError_t e = bin_rays<true>(s, w, &l);
                if (e != cudaSuccess) return e;
                wt.use_order = 1;
                if (ev) cudaEventRecord(ev[3], s);   // booked on the shading stage
            }
            k_shadow<<<cfg.grid_shadow, kBlock, cfg.smem_top, s>>>(sc, wt, WalkCfg{cfg.step_nodes, cfg.step_prims, cfg.min_active, cfg.node_min, cfg.prim_min});
        }
        l++;
    }
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
                         rtw_hit* hits, int32_t* occluded, uint32_t* work2, uint64_t* launches) {
    if (n <= 0) return cudaSuccess;
    int grid = (n + kBlock - 1) / kBlock;
    if (grid > cfg.grid_traverse) grid = cfg.grid_traverse;
    if (cfg.small) {
        k_trace_small<<<(n + kBlock - 1) / kBlock, kBlock, 0, s>>>(*cfg.small, rays, n, hits, occluded);
        if (launches) *launches += 1;
        return cudaGetLastError();
    }
    cudaError_t e = cudaMemsetAsync(work2, 0, 2 * sizeof(uint32_t), s);
    if (e != cudaSuccess) return e;
    k_trace<<<grid, kBlock, cfg.smem_top, s>>>(sc, rays, n, hits, occluded, work2, WalkCfg{cfg.step_nodes, cfg.step_prims, cfg.min_active, cfg.node_min, cfg.prim_min});
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
