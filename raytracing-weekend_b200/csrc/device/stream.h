// stream.h — per-thread bodies of the STREAMING wavefront (kernels/stream.cu).
//
// Same cut of the `rayColor` loop (raygen/raygen.cu:36-84) as wavefront.h, same per-path RNG
// ledger, same arithmetic in the same order — but a path is not a slot of a pool that kernels
// visit through index queues.  Its whole state (80 bytes, five float4) MOVES every iteration:
// the shading kernel reads it from the queue of the material class it hit, shades, traces the
// NEE ray and the next ray in the same thread, and appends the state to the queue of the class
// the next ray hit.  Reads and writes are consecutive float4 per lane (SoA), nothing is
// gathered, and the hit record, the radiance sum and the pixel travel with the path.
// Paths that end are flushed to the accumulator where they end (the terminal vertex — miss,
// light, normal material — is shaded by the thread that traced the ray, closehit.cu:63-67 via
// material/diffuseLight.cu:48-69, miss/miss.cu:23-30) and simply are not appended.
#pragma once
#include "wavefront.h"

namespace rtw {

// stream queues: RTW_SQ_CLASSES scattering classes (RTW_Q_LAMBERTIAN .. RTW_Q_ISOTROPIC), two
// buffers (read / append, swapped every iteration), RTW_SQ_FIELDS float4 arrays of P entries each
#define RTW_SQ_CLASSES 4
#define RTW_SQ_FIELDS  5

struct PathV {                 // one path in registers; field order = the five float4 of a queue entry
    float3 o;   float time;        // ray origin, ray time
    float3 d;   float gather;      // ray direction, gather time (PerRayData::gatherTime)
    float3 thr; uint32_t seed;     // throughput, path seed (PerRayData::seed)
    uint32_t depth, lseed;         // bounce count, ray-time seed (the local seed of rayColor)
    float t;    uint32_t leaf;     // closest hit of the ray: distance, leaf index
    float3 rad; uint32_t pixel;    // radiance sum of the sample, pixel index
};

RTW_HD float4* sq_field(const DWave& w, int buf, int cls, int field) {
    return w.sq + (size_t)((buf * RTW_SQ_CLASSES + cls) * RTW_SQ_FIELDS + field) * (size_t)w.P;
}
RTW_HD PathV sq_load(const DWave& w, int buf, int cls, uint32_t i) {
    const float4 a = sq_field(w, buf, cls, 0)[i];
    const float4 b = sq_field(w, buf, cls, 1)[i];
    const float4 c = sq_field(w, buf, cls, 2)[i];
    const float4 d = sq_field(w, buf, cls, 3)[i];
    const float4 e = sq_field(w, buf, cls, 4)[i];
    PathV p;
    p.o = f3(a); p.time = a.w;
    p.d = f3(b); p.gather = b.w;
    p.thr = f3(c); p.seed = f2u(c.w);
    p.depth = f2u(d.x); p.lseed = f2u(d.y); p.t = d.z; p.leaf = f2u(d.w);
    p.rad = f3(e); p.pixel = f2u(e.w);
    return p;
}
RTW_HD void sq_store(const DWave& w, int buf, int cls, uint32_t i, const PathV& p) {
    sq_field(w, buf, cls, 0)[i] = make_float4(p.o.x, p.o.y, p.o.z, p.time);
    sq_field(w, buf, cls, 1)[i] = make_float4(p.d.x, p.d.y, p.d.z, p.gather);
    sq_field(w, buf, cls, 2)[i] = make_float4(p.thr.x, p.thr.y, p.thr.z, u2f(p.seed));
    sq_field(w, buf, cls, 3)[i] = make_float4(u2f(p.depth), u2f(p.lseed), p.t, u2f(p.leaf));
    sq_field(w, buf, cls, 4)[i] = make_float4(p.rad.x, p.rad.y, p.rad.z, u2f(p.pixel));
}

RTW_HD PathV path_of(const NewPath& np) {
    PathV p;
    p.o = f3(np.ray.o); p.time = np.ray.o.w;
    p.d = f3(np.ray.d); p.gather = np.ray.d.w;
    p.thr = f3(np.st.thr); p.seed = f2u(np.st.thr.w);
    p.depth = misc_depth(np.st.misc.x); p.lseed = np.st.misc.y;
    p.t = 0.f; p.leaf = RTW_REF_NONE;
    p.rad = f3(np.rad); p.pixel = f2u(np.rad.w);
    return p;
}
// The closest-hit ray of a path (optixTraverse(tmin 1e-6, tmax RT_DEFAULT_MAX), raygen.cu:41-54).
RTW_HD RayW closest_ray(const PathV& p) {
    RayW r;
    r.o = p.o; r.d = p.d; r.tmin = 1e-6f; r.tmax = 1.e27f;
    r.time = p.time; r.gather = p.gather; r.seed = p.seed;
    return r;
}
// A ray for a lane that has nothing to trace while its warp walks: it costs the walk no slow path
// (no division by zero) and its result is dropped.
RTW_HD RayW idle_ray(float tmin, float tmax) {
    RayW r;
    r.o = f3(0.f); r.d = f3(1.f); r.tmin = tmin; r.tmax = tmax; r.time = 0.f; r.gather = 0.f; r.seed = 0u;
    return r;
}
RTW_HD void path_flush(const DWave& w, const PathV& p) {
    logic_flush(w, make_float4(p.rad.x, p.rad.y, p.rad.z, u2f(p.pixel)));
}

// One scattering vertex: __closesthit__radiance (shaders/closehit.cu:45-121) for a hit on a
// lambertian / metal / dielectric / isotropic primitive, then the tail of the rayColor loop body
// (raygen.cu:60-83).  Statement for statement the scattering half of shade_thread (wavefront.h),
// on registers.  Returns true iff the path goes on: then `p` holds the next ray and the updated
// state.  `sh` is the NEE ray (closehit.cu:70-118); its contribution is added by the caller iff
// the shadow ray is unoccluded.
// `kClass` is a RUN-TIME, warp-uniform value: one copy of the attribute / texture / NEE code serves the
// four classes (the instruction cache, 32 KB per SM, is what the fused kernels are short of).
template <class Mem>
RTW_HD bool shade_core(const DScene& sc, int max_depth, int kClass, PathV& p, ShadowOut& sh) {
    sh.emit = false;
    uint32_t seed = p.seed;
    uint32_t depth = p.depth;
    float3 throughput = p.thr;
    const float3 direction = p.d;
    RayW r = closest_ray(p);
    HitAttr h;
    hit_attributes<Mem>(sc, p.leaf, r, p.t, sc.has_images != 0, h);
    const float4 mat = ldg4(sc.mats + ldgi(sc.leaf_mat + p.leaf));
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
                sh.contrib = nee * throughput;
                sh.seed = seed;
            }
        }
    }

    // raygen.cu:60 (see shade_thread: the scattering materials return radiance 0, the sum changes
    // only when the product is not +-0)
    const float3 add = s.radiance * throughput;
    if (!(add.x == 0.0f && add.y == 0.0f && add.z == 0.0f)) {
        p.rad.x += add.x; p.rad.y += add.y; p.rad.z += add.z;
    }
    if (s.event != RTW_EV_HIT) return false;
    throughput = throughput * s.attenuation;
    if (2 <= depth) {  // raygen.cu:74-82 (Q8)
        const float probability = max3(throughput);
        if (probability < rnd(seed)) return false;
        throughput = throughput / probability;
    }
    depth++;
    if ((int)depth >= max_depth) return false;  // raygen.cu:36 (Q9)
    uint32_t lseed = p.lseed;
    p.time = rnd(lseed);  // raygen.cu:48
    p.o = s.origin;
    p.d = s.direction;
    p.thr = throughput;
    p.seed = seed;
    p.depth = depth;
    p.lseed = lseed;
    return true;
}

// What the closest hit of a path's ray means (closest_result + the RTW_Q_TERMINAL half of
// shade_thread): the scattering class whose queue the path joins, or -1 after the terminal vertex
// has been shaded and the sample flushed.
template <class Mem>
RTW_HD int classify_or_finish(const DScene& sc, const DWave& w, PathV& p, bool found, float t, uint32_t leaf) {
    float3 radiance = f3(0.f);
    if (!found) {
        radiance = miss_color(sc, p.d);
    } else {
        const int mt = meta_mattype(f2u(Mem::ld4(sc.prims + 2 * leaf + 1).w));
        p.t = t;
        p.leaf = leaf;
        if (mt == RTW_MAT_LAMBERTIAN) return RTW_Q_LAMBERTIAN;
        if (mt == RTW_MAT_METAL) return RTW_Q_METAL;
        if (mt == RTW_MAT_DIELECTRIC) return RTW_Q_DIELECTRIC;
        if (mt == RTW_MAT_ISOTROPIC) return RTW_Q_ISOTROPIC;
        const float4 mat = ldg4(sc.mats + ldgi(sc.leaf_mat + leaf));
        if ((int)f2u(mat.x) == RTW_MAT_DIFFUSE_LIGHT) {
            const RayW r = closest_ray(p);
            HitAttr h;
            hit_attributes<Mem>(sc, leaf, r, t, sc.has_images != 0, h);
            radiance = mat_diffuse_light(sc, (int)f2u(mat.y), p.d, h);
        }
        // RTW_MAT_NORMAL: Ray_Finish with radiance 0 (Q16)
    }
    const float3 add = radiance * p.thr;  // raygen.cu:60
    p.rad.x += add.x; p.rad.y += add.y; p.rad.z += add.z;
    path_flush(w, p);
    return -1;
}

}  // namespace rtw
