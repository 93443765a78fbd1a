// wavefront.h — per-thread bodies of the wavefront kernels.
//
// The reference runs one megakernel thread per pixel (raygen/raygen.cu:123-159:
// seed, jitter, camera ray, `rayColor` loop, gamma, store).  Here that loop is cut
// at every optixTraverse into stages over a pool of P path slots:
//   regenerate (flush finished paths, start new ones) → traverse closest →
//   shade (one queue per material) → traverse shadow → (next iteration)
// A path keeps the exact per-path RNG ledger of SURVEY.md §7b, so sample k of a
// pixel is the same random walk as in the scalar formulation.
//
// These functions are the whole per-thread work; the __global__ wrappers in
// kernels/wavefront.cu only add indexing, warp-aggregated queue pushes and
// shared-memory staging.
#pragma once
#include "material.h"
#include "traverse.h"

namespace rtw {

RTW_HD uint32_t misc_state(uint32_t m) { return (m >> 8) & 3u; }
RTW_HD uint32_t misc_depth(uint32_t m) { return m & 0xffu; }
RTW_HD uint32_t misc_pack(uint32_t depth, uint32_t state) { return (depth & 0xffu) | (state << 8); }

// ---- stage 0: flush a finished path ---------------------------------------
// removeNaNs + accumulation, raygen/raygen.cu:17-24,141.  On the device the add
// is an atomic (several slots may hold samples of one pixel).  `ra` is the slot's
// radiance record; a slot that never held a path carries RTW_PIXEL_NONE.
RTW_HD void logic_flush(const DWave& w, float4 ra) {
    if (f2u(ra.w) == RTW_PIXEL_NONE) return;
    if (!(ra.x == ra.x)) ra.x = 0.0f;
    if (!(ra.y == ra.y)) ra.y = 0.0f;
    if (!(ra.z == ra.z)) ra.z = 0.0f;
#if defined(__CUDA_ARCH__)
    // one 16-byte vector reduction (sm_90+) instead of four scalar ones: the four lanes of the
    // vector are independent fp32 adds, so the sums are the same bits
    atomicAdd(w.accum + f2u(ra.w), make_float4(ra.x, ra.y, ra.z, 1.0f));
#else
    float* a = reinterpret_cast<float*>(w.accum + f2u(ra.w));
    a[0] += ra.x; a[1] += ra.y; a[2] += ra.z; a[3] += 1.0f;
#endif
}
// A slot that is not refilled (no sample left to start): nothing to trace, nothing to flush again.
RTW_HD void retire_slot(const DWave& w, uint32_t slot) {
    RayRec none;
    none.o = make_float4(0.f, 0.f, 0.f, u2f(RTW_RAY_NONE));
    none.d = make_float4(0.f, 0.f, 0.f, 0.f);
    w.rec[slot].ray = none;
    PathRec st;
    st.thr = make_float4(0.f, 0.f, 0.f, 0.f);
    st.misc = make_uint2(misc_pack(0, RTW_ST_EMPTY), 0u);
    st.hit = make_uint2(0u, 0u);
    w.rec[slot].st = st;
    w.rad[slot] = make_float4(0.f, 0.f, 0.f, u2f(RTW_PIXEL_NONE));
}

// ---- stage 0b: start sample `ordinal` in a free slot -----------------------
// raygen.cu:129-135 (seed, jitter), shaders/camera.cu:11-19 (camera ray, Q3),
// raygen.cu:89-95 (color(): thePrd.seed = S4, gatherTime from the local seed, Q4),
// raygen.cu:30,48 (rayColor(): local seed copy, first ray time).
struct NewPath {
    RayRec ray;
    PathRec st;
    float4 rad;
};
RTW_HD NewPath make_path(const DScene& sc, const DWave& w, unsigned long long ordinal) {
    const unsigned long long npix = (unsigned long long)w.pixel_count;     // the tile of this call (the whole image: Nx * Ny)
    const uint32_t pixel = w.pixel_first + (uint32_t)(ordinal % npix);
    const uint32_t ki = (uint32_t)(ordinal / npix);
    const uint32_t k = (uint32_t)w.sample_first + ki * (uint32_t)w.sample_stride;
    const uint32_t x = pixel % (uint32_t)w.Nx, y = pixel / (uint32_t)w.Nx;
    uint32_t seed = tea<64>(pixel, k);  // Q1: k == 0 is the reference's only sample
    const float s = float((float)x + rnd(seed)) / float((uint32_t)w.Nx);
    const float t = float((float)y + rnd(seed)) / float((uint32_t)w.Ny);
    const float3 rd = sc.cam_lens_radius * random_in_unit_disk(seed);
    const float3 lens_offset = sc.cam_u * rd.x + sc.cam_v * rd.y;
    const float3 origin = sc.cam_origin + lens_offset;
    const float3 direction = sc.cam_ll + (s * sc.cam_horizontal) + (t * sc.cam_vertical) - origin;
    const uint32_t s4 = seed;
    const float gather = sc.cam_time0 + rnd(seed) * (sc.cam_time1 - sc.cam_time0);
    uint32_t lseed = s4;
    const float ray_time = rnd(lseed);
    NewPath np;
    np.ray.o = make_float4(origin.x, origin.y, origin.z, ray_time);
    np.ray.d = make_float4(direction.x, direction.y, direction.z, gather);
    np.st.thr = make_float4(1.0f, 1.0f, 1.0f, u2f(s4));
    np.st.hit = make_uint2(0u, 0u);
    np.rad = make_float4(0.f, 0.f, 0.f, u2f(pixel));
    // `while (depth < maxRayDepth)` (raygen.cu:36): with a depth limit of 0 no ray is traced
    np.st.misc = make_uint2(misc_pack(0, w.max_depth > 0 ? RTW_ST_ALIVE : RTW_ST_DEAD), lseed);
    return np;
}
RTW_HD void store_path(const DWave& w, uint32_t slot, const NewPath& np) {
    w.rec[slot].ray = np.ray;
    w.rec[slot].st = np.st;
    w.rad[slot] = np.rad;
}
RTW_HD void generate_path(const DScene& sc, const DWave& w, uint32_t slot, unsigned long long ordinal) {
    store_path(w, slot, make_path(sc, w, ordinal));
}

// ---- stage 1: closest-hit traversal + material classification --------------
// optixTraverse(tmin 1e-6, tmax RT_DEFAULT_MAX), raygen.cu:41-54.  Returns the
// shading queue class of the slot, or -1 if the slot holds no live path.
// How a stage finds its hits: the LBVH (global memory, top levels staged by the kernel) or the
// by-value copy of a small scene.
template <class Mem>
struct TravBvh {
    const DScene& sc;
    const float4* top;
    RTW_HD bool closest(const RayW& r, float& t, uint32_t& leaf) const { return traverse_closest<Mem, false>(sc, top, r, t, leaf, nullptr); }
    RTW_HD bool any(const RayW& r) const { return traverse_any<Mem, false>(sc, top, r, nullptr); }
};
struct TravSmall {
    const SmallScene& ss;
    RTW_HD bool closest(const RayW& r, float& t, uint32_t& leaf) const { return small_closest(ss, r, t, leaf); }
    RTW_HD bool any(const RayW& r) const { return small_any(ss, r); }
};

// The closest-hit ray of a slot (false: the slot holds no live ray), and what becomes of its result.
RTW_HD bool closest_ray_of(const DScene& sc, const DWave& w, uint32_t slot, RayW& r) {
    const RayRec ray = w.rec[slot].ray;
    const float4 ro = ray.o, rd = ray.d;
    r.o = f3(ro);
    r.d = f3(rd);
    r.tmin = 1e-6f;
    r.tmax = 1.e27f;  // shaders/appconfig.h:3
    r.time = ro.w;
    r.gather = rd.w;
    r.seed = 0u;
    if (f2u(ro.w) == RTW_RAY_NONE) return false;
    if (sc.has_media) r.seed = f2u(w.rec[slot].st.thr.w);
    return true;
}
template <class Mem>
RTW_HD int closest_result(const DScene& sc, const DWave& w, uint32_t slot, bool found, float t, uint32_t leaf) {
    // the hit rides in the last 8 bytes of the slot record: the shading gather then touches ONE 64-byte line per slot
    // (profiles/r02/hit_in_record_r02g.md: k_shade_all 396 -> 344 MB of DRAM traffic per launch, C2 +1.8 %)
    w.rec[slot].st.hit = make_uint2(f2u(t), found ? leaf : RTW_REF_NONE);
    if (!found) return RTW_Q_TERMINAL;
    const int mt = meta_mattype(f2u(Mem::ld4(sc.prims + 2 * leaf + 1).w));
    if (mt == RTW_MAT_LAMBERTIAN) return RTW_Q_LAMBERTIAN;
    if (mt == RTW_MAT_METAL) return RTW_Q_METAL;
    if (mt == RTW_MAT_DIELECTRIC) return RTW_Q_DIELECTRIC;
    if (mt == RTW_MAT_ISOTROPIC) return RTW_Q_ISOTROPIC;
    return RTW_Q_TERMINAL;  // diffuse light, normal material
}
template <class Mem, class Trav>
RTW_HD int traverse_thread_with(const DScene& sc, const Trav& trav, const DWave& w, uint32_t slot) {
    RayW r;
    if (!closest_ray_of(sc, w, slot, r)) return -1;
    float t;
    uint32_t leaf;
    const bool found = trav.closest(r, t, leaf);
    return closest_result<Mem>(sc, w, slot, found, t, leaf);
}
template <class Mem>
RTW_HD int traverse_thread(const DScene& sc, const float4* top, const DWave& w, uint32_t slot) {
    return traverse_thread_with<Mem>(sc, TravBvh<Mem>{sc, top}, w, slot);
}

// ---- stage 2: shading -------------------------------------------------------
struct ShadowOut {
    bool emit;
    float3 o, d;
    float tmax;
    float3 contrib;
    uint32_t seed;
    float gather;     // PerRayData::gatherTime of the path (moving spheres)
};

// __closesthit__radiance (shaders/closehit.cu:45-121) / __miss__Program
// (miss/miss.cu:23-30) followed by the tail of the rayColor loop body
// (raygen.cu:60-83: accumulate, continue/terminate, Russian roulette, depth++).
// Returns true iff the path ended here (the slot goes to the regen queue).
// `ahead`: a slot this thread will shade later (or RTW_REF_NONE).  Its record and hit are requested
// into L2 right after this slot's own loads are issued, so the DRAM latency of the gather through
// the queue index is paid while this vertex is shaded.
RTW_HD void prefetch_slot(const DWave& w, uint32_t ahead) {
#if defined(__CUDA_ARCH__)
    if (ahead != RTW_REF_NONE) {
        asm volatile("prefetch.global.L2 [%0];" ::"l"(w.rec + ahead));
    }
#else
    (void)w; (void)ahead;
#endif
}
template <int kClass, class Mem>
RTW_HD bool shade_thread(const DScene& sc, const DWave& w, uint32_t slot, ShadowOut& sh, uint32_t ahead = RTW_REF_NONE) {
    sh.emit = false;
    const RayRec ray = w.rec[slot].ray;
    const PathRec st = w.rec[slot].st;
    const uint2 hit = st.hit;
    prefetch_slot(w, ahead);
    const float4 ro = ray.o, rd = ray.d;
    const float4 th = st.thr;
    float4 ra = make_float4(0.f, 0.f, 0.f, 0.f);
    if (kClass == RTW_Q_TERMINAL) ra = w.rad[slot];   // the scattering classes touch the sum only in the rare case below
    const uint2 m = st.misc;
    uint32_t depth = misc_depth(m.x);
    uint32_t lseed = m.y;
    uint32_t seed = f2u(th.w);
    float3 throughput = f3(th);
    const float3 direction = f3(rd);
    const uint32_t leaf = hit.y;
    const float t = u2f(hit.x);

    if (kClass == RTW_Q_TERMINAL) {
        float3 radiance = f3(0.f);
        if (leaf == RTW_REF_NONE) {
            radiance = miss_color(sc, direction);
        } else {
            const float4 mat = ldg4(sc.mats + ldgi(sc.leaf_mat + leaf));
            if ((int)f2u(mat.x) == RTW_MAT_DIFFUSE_LIGHT) {
                RayW r;
                r.o = f3(ro); r.d = direction; r.tmin = 1e-6f; r.tmax = 1.e27f;
                r.time = ro.w; r.gather = rd.w; r.seed = seed;
                HitAttr h;
                hit_attributes<Mem>(sc, leaf, r, t, sc.has_images != 0, h);
                radiance = mat_diffuse_light(sc, (int)f2u(mat.y), direction, h);
            }
            // RTW_MAT_NORMAL: Ray_Finish with radiance 0 (Q16)
        }
        const float3 add = radiance * throughput;  // raygen.cu:60
        ra.x += add.x; ra.y += add.y; ra.z += add.z;
        w.rad[slot] = ra;   // the path record is not needed again: the slot goes to the regen queue
        return true;
    }

    RayW r;
    r.o = f3(ro); r.d = direction; r.tmin = 1e-6f; r.tmax = 1.e27f;
    r.time = ro.w; r.gather = rd.w; r.seed = seed;
    HitAttr h;
    hit_attributes<Mem>(sc, leaf, r, t, sc.has_images != 0, h);
    const float4 mat = ldg4(sc.mats + ldgi(sc.leaf_mat + leaf));
    const int tex = (int)f2u(mat.y);
    Scatter s;
    if (kClass == RTW_Q_LAMBERTIAN) mat_lambertian(sc, tex, h, seed, s);
    else if (kClass == RTW_Q_METAL) mat_metal(sc, tex, mat.z, direction, h, seed, s);
    else if (kClass == RTW_Q_DIELECTRIC) mat_dielectric(mat.z, direction, h, seed, s);
    else mat_isotropic(sc, tex, h, seed, s);

    // next-event estimation, closehit.cu:70-118 (only lambertian is non-specular, Q15)
    if (kClass == RTW_Q_LAMBERTIAN && s.event == RTW_EV_HIT && sc.n_lights > 0 && sc.has_light_rect) {
        int indexLight = 0;
        if (1 < sc.n_lights) {
            int v = (int)floorf(rnd(seed) * sc.n_lights);
            indexLight = v < 0 ? 0 : (v > sc.n_lights - 1 ? sc.n_lights - 1 : v);
        }
        const LightSample ls = rect_generate(sc, indexLight, s.origin, seed);
        if (ls.pdf > 0) {
            const float3 f = s.attenuation * RTW_INV_PI;  // eval_bsdf_diffuse_reflection
            const float bpdf = fmaxf(0.0f, dot(ls.direction, h.normal) * RTW_INV_PI);
            if (0.0f < bpdf && (f.x != 0.0f || f.y != 0.0f || f.z != 0.0f)) {
                const float weight = powerHeuristic(ls.pdf, bpdf);
                const float3 nee = f * ls.emission * (weight * dot(ls.direction, h.normal) / ls.pdf);
                sh.emit = true;
                sh.o = s.origin;
                sh.d = ls.direction;
                sh.tmax = ls.distance - 500 * 1.0e-7f;
                sh.contrib = nee * throughput;  // added by the shadow stage iff unoccluded
                sh.seed = seed;
                sh.gather = rd.w;
            }
        }
    }

    // raygen.cu:60: sampleRadiance += radiance * throughput.  The scattering materials return radiance 0
    // (NEE arrives through the shadow stage), so the sum only changes when the product is not +-0 —
    // i.e. when throughput is inf/NaN; x + (+-0) == x for every x the sum can hold (it is never -0).
    const float3 add = s.radiance * throughput;
    const bool touch_rad = !(add.x == 0.0f && add.y == 0.0f && add.z == 0.0f);
    uint32_t state = RTW_ST_ALIVE;
    float3 origin = f3(ro);
    float3 dir = direction;
    if (s.event != RTW_EV_HIT) {
        state = RTW_ST_DEAD;
    } else {
        origin = s.origin;
        dir = s.direction;
        throughput = throughput * s.attenuation;
        if (2 <= depth) {  // raygen.cu:74-82 (Q8)
            const float probability = max3(throughput);
            if (probability < rnd(seed)) state = RTW_ST_DEAD;
            else throughput = throughput / probability;
        }
        if (state == RTW_ST_ALIVE) {
            depth++;
            if ((int)depth >= w.max_depth) state = RTW_ST_DEAD;  // raygen.cu:36 (Q9)
        }
    }
    if (state == RTW_ST_ALIVE) {
        const float ray_time = rnd(lseed);  // raygen.cu:48
        RayRec next;
        next.o = make_float4(origin.x, origin.y, origin.z, ray_time);
        next.d = make_float4(dir.x, dir.y, dir.z, rd.w);
        w.rec[slot].ray = next;
        PathRec ns;
        ns.thr = make_float4(throughput.x, throughput.y, throughput.z, u2f(seed));
        ns.misc = make_uint2(misc_pack(depth, state), lseed);
        ns.hit = make_uint2(0u, 0u);
        w.rec[slot].st = ns;
    }
    if (touch_rad) {
        float4 ra = w.rad[slot];
        ra.x += add.x; ra.y += add.y; ra.z += add.z;
        w.rad[slot] = ra;
    }
    return state != RTW_ST_ALIVE;
}

// ---- stage 3: shadow rays ----------------------------------------------------
// traceOcclusion, shaders/closehit.cu:16-42,97-113: tmin 500e-7, ray time 0.
RTW_HD void shadow_ray_of(const DScene& sc, const DWave& w, uint32_t i, RayW& r, float4& c, uint32_t& slot) {
    const float4 so = w.sh_o[i];
    const float4 sd = w.sh_d[i];
    c = w.sh_c[i];
    slot = f2u(sd.w);
    r.o = f3(so);
    r.d = f3(sd);
    r.tmin = 500 * 1.0e-7f;
    r.tmax = so.w;
    r.time = 0.0f;
    r.gather = sc.has_motion ? w.rec[slot].ray.d.w : 0.f;
    r.seed = f2u(c.w);
}
RTW_HD void shadow_result(const DWave& w, uint32_t slot, float4 c, bool occluded) {
    if (!occluded) {
        float4 ra = w.rad[slot];
        ra.x += c.x; ra.y += c.y; ra.z += c.z;
        w.rad[slot] = ra;
    }
}
template <class Trav>
RTW_HD void shadow_thread_with(const DScene& sc, const Trav& trav, const DWave& w, uint32_t i) {
    RayW r;
    float4 c;
    uint32_t slot;
    shadow_ray_of(sc, w, i, r, c, slot);
    shadow_result(w, slot, c, trav.any(r));
}
template <class Mem>
RTW_HD void shadow_thread(const DScene& sc, const float4* top, const DWave& w, uint32_t i) {
    shadow_thread_with(sc, TravBvh<Mem>{sc, top}, w, i);
}

}  // namespace rtw
