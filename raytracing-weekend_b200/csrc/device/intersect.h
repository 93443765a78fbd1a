// intersect.h — primitive tests ("candidate" functions) and hit attributes.
//
// Restates the reference's intersection programs for a single ray:
//   geometry/sphere.cu:48-120          sphere
//   geometry/movingSphere.cu:33-39,61-135   moving sphere (Q24 literal)
//   shaders/aarectx.cu:8-49, aarecty.cu:8-50, aarectz.cu:9-51   axis-aligned rects
//   geometry/volumeBox.cu:29-113       constant medium in a box   (Q23)
//   geometry/volumeSphere.cu:37-127    constant medium in a sphere (Q23)
// OptiX runtime semantics supplied here: object ray = world→object * world ray,
// optixReportIntersection accepts iff tmin <= t <= tmax, and the closest hit is
// the lexicographic minimum of (t, primitive index) — independent of traversal order.
#pragma once
#include "detmath.h"
#include "layout.h"
#include "rng.h"

namespace rtw {

// memory policies: read-only global path vs generic (shared-memory staged scene)
struct MemGlobal {
    static RTW_HD float4 ld4(const float4* p) {
#if defined(__CUDA_ARCH__)
        return __ldg(p);
#else
        return *p;
#endif
    }
};
struct MemGeneric {
    static RTW_HD float4 ld4(const float4* p) { return *p; }
};

struct RayW {
    float3 o, d;
    float tmin, tmax;
    float time;     // ray time: motion-transform key interpolation
    float gather;   // PerRayData::gatherTime: moving-sphere centre
    uint32_t seed;  // path seed at ray start (constant media, Q23)
};

// Matrix-motion transform of the moving sphere's instance child
// (geometry/ioMovingSphere.h:162-182): keys translate(c0), translate(c1) over [0,1].
RTW_HD float3 motion_translation(float4 a, float4 b, float ray_time) {
    float f = ray_time;
    if (f < 0.f) f = 0.f;
    if (f > 1.f) f = 1.f;
    return f3(a) * (1.0f - f) + f3(b) * f;
}
// geometry/movingSphere.cu:33-39
RTW_HD float3 moving_center(float4 a, float4 b, float2 times, float time) {
    if (times.x == times.y) return f3(a);
    return f3(a) + ((time - times.x) / (times.y - times.x)) * (f3(b) - f3(a));
}

// geometry/volumeBox.cu:29-52
RTW_HD bool hit_boundary_box(float3 direction, float3 origin, float3 boxMin, float3 boxMax, float tMin, float tMax, float& rec) {
    float3 t0 = (boxMin - origin) / direction;
    float3 t1 = (boxMax - origin) / direction;
    float temp1 = max3(vmin(t0, t1));
    float temp2 = min3(vmax(t0, t1));
    if (temp1 > temp2) return false;
    if (temp1 < tMax && temp1 > tMin) { rec = temp1; return true; }
    if (temp2 < tMax && temp2 > tMin) { rec = temp2; return true; }
    return false;
}
// geometry/volumeSphere.cu:37-61
RTW_HD bool hit_boundary_sphere(float3 direction, float3 origin, float3 center, float radius, float tmin, float tmax, float& rec) {
    float3 oc = origin - center;
    float a = dot(direction, direction);
    float b = dot(oc, direction);
    float c = dot(oc, oc) - radius * radius;
    float discriminant = b * b - a * c;
    if (discriminant < 0.f) return false;
    float t = (-b - sqrtf(discriminant)) / a;
    if (t < tmax && t > tmin) { rec = t; return true; }
    t = (-b + sqrtf(discriminant)) / a;
    if (t < tmax && t > tmin) { rec = t; return true; }
    return false;
}

// Where the traversal reads a leaf from.  SrcScene: the packed arrays in global memory
// (per-lane leaf index).  SrcSmall (traverse.h): the kernel-parameter copy of a scene of
// <= RTW_SMALL_MAX leaves, indexed by a warp-uniform loop counter.
template <class Mem>
struct SrcScene {
    const DScene& sc;
    RTW_HD void prim(uint32_t leaf, float4& a, float4& b) const {
        a = Mem::ld4(sc.prims + 2 * leaf);
        b = Mem::ld4(sc.prims + 2 * leaf + 1);
    }
    RTW_HD void w2o(int xf, float4& r0, float4& r1, float4& r2) const {
        const float4* m = sc.xforms + 6 * xf;
        r0 = Mem::ld4(m); r1 = Mem::ld4(m + 1); r2 = Mem::ld4(m + 2);
    }
    RTW_HD float2 ext(uint32_t leaf) const {
#if defined(__CUDA_ARCH__)
        return __ldg(sc.prim_ext + leaf);
#else
        return sc.prim_ext[leaf];
#endif
    }
    RTW_HD int prim_id(uint32_t leaf) const {
#if defined(__CUDA_ARCH__)
        return __ldg(sc.leaf_prim + leaf);
#else
        return sc.leaf_prim[leaf];
#endif
    }
};
struct SrcSmall {
    const SmallScene& ss;
    RTW_HD void prim(uint32_t leaf, float4& a, float4& b) const { a = ss.leaf[leaf].a; b = ss.leaf[leaf].b; }
    RTW_HD void w2o(int xf, float4& r0, float4& r1, float4& r2) const {
        r0 = ss.w2o[3 * xf]; r1 = ss.w2o[3 * xf + 1]; r2 = ss.w2o[3 * xf + 2];
    }
    RTW_HD float2 ext(uint32_t leaf) const { return ss.leaf[leaf].ext; }
    RTW_HD int prim_id(uint32_t leaf) const { return ss.leaf[leaf].prim_id; }
};

// The world ray carried through an instance transform, remembered between the
// primitives of one traversal: consecutive leaves mostly share their instance (the six
// faces of a box, the 1000 small spheres of scene 4), so the 3x4 product is done once
// per run of equal transform ids instead of once per primitive.  Same values either way.
struct XfRay {
    int xf;        // transform id of (o, d); -1 = nothing cached
    float3 o, d;
};

// ---- the candidate tests, one per primitive kind, on an object-space ray ----
// shaders/aarectx.cu:8-49 (aarecty, aarectz alike): plane distance from the normal axis
// (on, dn), in-plane coordinates from the other two.
RTW_HD bool rect_candidate(float4 a, float k, float on, float dn, float o1, float d1, float o2, float d2, float tmin,
                           float tmax, float& t_out) {
    // The reference: t = (k - on) / dn; reject unless tmin <= t <= tmax (NaN rejected) and the point lies
    // in [a.x, a.y] x [a.z, a.w] (comparisons with a NaN coordinate pass).  Written without early exits:
    // a warp walks a leaf range in lockstep and rarely has all 32 lanes failing the same test.
    //
    // A path leaving a rect has its origin ON that plane about half of the time, so k - on is exactly 0
    // for one primitive per ray; 0 / dn = +-0 (NaN when dn is 0 or NaN) would take the division's
    // out-of-line slow path on a few lanes in every iteration.  Neither +-0 nor NaN lies in [tmin, tmax]
    // when tmin > 0 (every ray of the path; only test-hook rays may carry tmin <= 0), so the quotient is
    // not needed then: the lane divides 1 by dn instead and drops the result.
    const float num = k - on;
    const bool on_plane = num == 0.0f;
    if (on_plane && !(tmin > 0.0f)) {
        const float t = num * ((dn > 0.0f || dn < 0.0f) ? copysignf(1.0f, dn) : NAN);
        const float pa = o1 + t * d1, pb = o2 + t * d2;
        t_out = t;
        return (t >= tmin) & (t <= tmax) & !((pa < a.x) | (pa > a.y) | (pb < a.z) | (pb > a.w));
    }
    const float t = (on_plane ? 1.0f : num) / dn;
    const float pa = o1 + t * d1;
    const float pb = o2 + t * d2;
    t_out = t;
    return !on_plane & (t >= tmin) & (t <= tmax) & !((pa < a.x) | (pa > a.y) | (pb < a.z) | (pb > a.w));
}
// geometry/sphere.cu:48-120, geometry/movingSphere.cu:61-135 (centre c already resolved)
RTW_HD bool sphere_candidate(float3 c, float radius, float3 o, float3 d, float tmin, float tmax, float& t_out) {
    float3 oc = o - c;
    float qa = dot(d, d);
    float qb = dot(oc, d);
    float qc = dot(oc, oc) - radius * radius;
    float discriminant = qb * qb - qa * qc;
    if (discriminant < 0.0f) return false;
    float sq = sqrtf(discriminant);
    float t = (-qb - sq) / qa;
    if (t < tmax && t > tmin) { t_out = t; return true; }
    t = (-qb + sq) / qa;
    if (t < tmax && t > tmin) { t_out = t; return true; }
    return false;
}
// geometry/volumeBox.cu:54-113, geometry/volumeSphere.cu:63-127; EPSILON is `constexpr int` == 0
// in the reference (Q23); the free-flight draw is keyed by (ray seed, primitive id).
RTW_HD bool medium_candidate(bool box, float4 a, float4 b, int prim, float3 o, float3 d, const RayW& r, float& t_out) {
    const float tmin = r.tmin, tmax = r.tmax;
    float hitt1, hitt2, density;
    if (box) {
        density = a.w;
        float3 bmin = f3(a), bmax = f3(b);
        if (!hit_boundary_box(d, o, bmin, bmax, -RTW_FLT_MAX, RTW_FLT_MAX, hitt1)) return false;
        if (!hit_boundary_box(d, o, bmin, bmax, hitt1 + 0, RTW_FLT_MAX, hitt2)) return false;
    } else {
        density = b.x;
        float3 c = f3(a);
        if (!hit_boundary_sphere(d, o, c, a.w, -RTW_FLT_MAX, RTW_FLT_MAX, hitt1)) return false;
        if (!hit_boundary_sphere(d, o, c, a.w, hitt1 + 0, RTW_FLT_MAX, hitt2)) return false;
    }
    if (hitt1 < tmin) hitt1 = tmin;
    if (hitt2 > tmax) hitt2 = tmax;
    if (hitt1 >= hitt2) return false;
    if (hitt1 < 0.f) hitt1 = 0.f;
    uint32_t st = medium_state(r.seed, (uint32_t)prim);
    float u = box ? randf(st) : rnd(st);  // volumeBox.cu:79 / volumeSphere.cu:93
    float hitDistance = -(1.f / density) * det_logf(u);
    float hitt = hitt1 + (hitDistance / length(d));
    if (!(hitt >= tmin && hitt <= tmax)) return false;
    t_out = hitt;
    return true;
}

// Object-space ray of the primitive (a, b, meta).
template <class Src>
RTW_HD void object_ray(const Src& src, float4 a, float4 b, uint32_t meta, const RayW& r, XfRay& cache, float3& o,
                       float3& d) {
    const int type = meta_type(meta);
    o = r.o;
    d = r.d;
    if (type == RTW_PRIM_MOVING_SPHERE) {
        o = r.o - motion_translation(a, b, r.time);
    } else {
        const int xf = meta_xform(meta);
        if (xf >= 0) {
            if (xf != cache.xf) {
                float4 r0, r1, r2;
                src.w2o(xf, r0, r1, r2);
                cache.o = xform_point(r0, r1, r2, r.o);
                cache.d = xform_vector(r0, r1, r2, r.d);
                cache.xf = xf;
            }
            o = cache.o;
            d = cache.d;
        }
    }
}

// True (and the accepted distance) iff leaf `leaf` reports an intersection that
// optixReportIntersection would accept for ray r.  Lanes of a warp may hold different
// leaves here: the three rect axes are one instruction stream (selects, no branches).
template <class Src>
RTW_HD bool prim_candidate(const Src& src, uint32_t leaf, const RayW& r, XfRay& cache, float& t_out) {
    float4 a, b;
    src.prim(leaf, a, b);
    const uint32_t meta = f2u(b.w);
    const int type = meta_type(meta);
    float3 o, d;
    object_ray(src, a, b, meta, r, cache, o, d);
    if (type == RTW_PRIM_RECT_X || type == RTW_PRIM_RECT_Y || type == RTW_PRIM_RECT_Z) {
        //   X: t from x, (pa, pb) = (y, z);  Y: t from y, (x, z);  Z: t from z, (x, y)
        const bool ax = type == RTW_PRIM_RECT_X, az = type == RTW_PRIM_RECT_Z;
        const float on = ax ? o.x : (az ? o.z : o.y), dn = ax ? d.x : (az ? d.z : d.y);
        const float o1 = ax ? o.y : o.x, d1 = ax ? d.y : d.x;
        const float o2 = az ? o.y : o.z, d2 = az ? d.y : d.z;
        return rect_candidate(a, b.x, on, dn, o1, d1, o2, d2, r.tmin, r.tmax, t_out);
    }
    if (type == RTW_PRIM_SPHERE || type == RTW_PRIM_MOVING_SPHERE) {
        float3 c = f3(a);
        if (type == RTW_PRIM_MOVING_SPHERE) c = moving_center(a, b, src.ext(leaf), r.gather);
        return sphere_candidate(c, a.w, o, d, r.tmin, r.tmax, t_out);
    }
    return medium_candidate(type == RTW_PRIM_VOLUME_BOX, a, b, src.prim_id(leaf), o, d, r, t_out);
}

// The same function for a leaf index that is UNIFORM over the warp (every lane walks the one
// leaf range of a small scene in lockstep): the kind and the transform id of the leaf are then
// warp-uniform, so the dispatch is by branch (no selects, no per-lane decode), and `cur` always
// holds the ray to test with: the world ray (cur.xf == -1) or its image under transform cur.xf.
template <class Src>
RTW_HD void retarget(const Src& src, int xf, const RayW& r, XfRay& cur) {
    if (xf == cur.xf) return;
    cur.xf = xf;
    if (xf >= 0) {
        float4 r0, r1, r2;
        src.w2o(xf, r0, r1, r2);
        cur.o = xform_point(r0, r1, r2, r.o);
        cur.d = xform_vector(r0, r1, r2, r.d);
    } else {
        cur.o = r.o;
        cur.d = r.d;
    }
}
template <class Src>
RTW_HD bool prim_candidate_uniform(const Src& src, uint32_t leaf, const RayW& r, XfRay& cur, float& t_out) {
    float4 a, b;
    src.prim(leaf, a, b);
    const uint32_t meta = f2u(b.w);
    const int type = meta_type(meta);
    if (type != RTW_PRIM_MOVING_SPHERE) {
        retarget(src, meta_xform(meta), r, cur);
        const float3 o = cur.o, d = cur.d;
        if (type == RTW_PRIM_RECT_X) return rect_candidate(a, b.x, o.x, d.x, o.y, d.y, o.z, d.z, r.tmin, r.tmax, t_out);
        if (type == RTW_PRIM_RECT_Y) return rect_candidate(a, b.x, o.y, d.y, o.x, d.x, o.z, d.z, r.tmin, r.tmax, t_out);
        if (type == RTW_PRIM_RECT_Z) return rect_candidate(a, b.x, o.z, d.z, o.x, d.x, o.y, d.y, r.tmin, r.tmax, t_out);
        if (type == RTW_PRIM_SPHERE) return sphere_candidate(f3(a), a.w, o, d, r.tmin, r.tmax, t_out);
        return medium_candidate(type == RTW_PRIM_VOLUME_BOX, a, b, src.prim_id(leaf), o, d, r, t_out);
    }
    const float3 o = r.o - motion_translation(a, b, r.time);
    return sphere_candidate(moving_center(a, b, src.ext(leaf), r.gather), a.w, o, r.d, r.tmin, r.tmax, t_out);
}

struct HitAttr {    // HitRecord, lib/raydata.cuh:117-124
    float3 point;
    float3 normal;
    float u, v;
};

// geometry/sphere.cu:24-32
RTW_HD void get_sphere_uv(float3 p, float& u, float& v) {
    float phi = det_atan2f(p.z, p.x);
    float theta = det_asinf(p.y);
    u = 1 - (phi + RTW_PI) / (2.f * RTW_PI);
    v = (theta + RTW_PIO2) / RTW_PI;
}

// Attributes of the winning intersection program (normal, uv, world point).
template <class Mem>
RTW_HD void hit_attributes(const DScene& sc, uint32_t leaf, const RayW& r, float t, bool want_uv, HitAttr& h) {
    const float4 a = Mem::ld4(sc.prims + 2 * leaf);
    const float4 b = Mem::ld4(sc.prims + 2 * leaf + 1);
    const uint32_t meta = f2u(b.w);
    const int type = meta_type(meta);
    const int xf = (type == RTW_PRIM_MOVING_SPHERE) ? -1 : meta_xform(meta);
    // The transform rows are loaded where they are used (and the world→object rows a second time for the
    // normal, an L1 hit) instead of living in twelve registers across the whole function: untransformed
    // primitives are the common case and the shading kernel is register-bound.
    const float4* m = sc.xforms + 6 * (xf >= 0 ? xf : 0);
    float3 o = r.o, d = r.d;   // object_ray()
    if (type == RTW_PRIM_MOVING_SPHERE) {
        o = r.o - motion_translation(a, b, r.time);
    } else if (xf >= 0) {
        const float4 w0 = Mem::ld4(m), w1 = Mem::ld4(m + 1), w2 = Mem::ld4(m + 2);
        o = xform_point(w0, w1, w2, r.o);
        d = xform_vector(w0, w1, w2, r.d);
    }
    float3 pobj = o + t * d;
    float3 point = pobj;
    if (type == RTW_PRIM_MOVING_SPHERE) {
        point = pobj + motion_translation(a, b, r.time);
    } else if (xf >= 0) {
        const float4 q0 = Mem::ld4(m + 3), q1 = Mem::ld4(m + 4), q2 = Mem::ld4(m + 5);
        point = xform_point(q0, q1, q2, pobj);
    }
    auto normal_to_world = [&](float3 n) -> float3 {
        if (xf < 0) return n;
        const float4 w0 = Mem::ld4(m), w1 = Mem::ld4(m + 1), w2 = Mem::ld4(m + 2);
        return xform_normal(w0, w1, w2, n);
    };
    h.point = point;
    h.u = 0.f;
    h.v = 0.f;
    if (type == RTW_PRIM_SPHERE || type == RTW_PRIM_MOVING_SPHERE) {
        // Q21: world-space point minus OBJECT-space centre, not normalised
        float3 c = f3(a);
        if (type == RTW_PRIM_MOVING_SPHERE) {
#if defined(__CUDA_ARCH__)
            float2 times = __ldg(sc.prim_ext + leaf);
#else
            float2 times = sc.prim_ext[leaf];
#endif
            c = moving_center(a, b, times, r.gather);
        }
        float3 n = (point - c) / a.w;
        h.normal = normal_to_world(n);
        // geometry/sphere.cu:66-69: get_sphere_uv takes the WORLD-space shading normal
        if (want_uv) get_sphere_uv(h.normal, h.u, h.v);
    } else if (type == RTW_PRIM_RECT_X || type == RTW_PRIM_RECT_Y || type == RTW_PRIM_RECT_Z) {
        float pa, pb;
        float3 n;
        if (type == RTW_PRIM_RECT_X) { pa = o.y + t * d.y; pb = o.z + t * d.z; n = f3(1.f, 0.f, 0.f); }
        else if (type == RTW_PRIM_RECT_Y) { pa = o.x + t * d.x; pb = o.z + t * d.z; n = f3(0.f, 1.f, 0.f); }
        else { pa = o.x + t * d.x; pb = o.y + t * d.y; n = f3(0.f, 0.f, 1.f); }
        if (meta_flip(meta)) n = -n;
        h.normal = normalize(normal_to_world(n));
        if (want_uv) {   // (u, v) only feed image textures (texture/imageTexture.cu:15); checker and noise use the point
            h.u = (pa - a.x) / (a.y - a.x);
            h.v = (pb - a.z) / (a.w - a.z);
        }
    } else {  // volumes: normal (1,0,0), uv 0 (volumeBox.cu:88-93, volumeSphere.cu:99-105)
        float3 n = f3(1.f, 0.f, 0.f);
        h.normal = normalize(normal_to_world(n));
    }
}

}  // namespace rtw
