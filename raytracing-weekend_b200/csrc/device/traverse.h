// traverse.h — stack-based traversal of the two-child LBVH.
//
// Replaces the driver-internal optixTraverse of the reference
// (raygen/raygen.cu:41-54 closest hit; shaders/closehit.cu:27-40 any hit with
// OPTIX_RAY_FLAG_TERMINATE_ON_FIRST_HIT).  Node = 64 B = 4 float4:
//   n0 = lmin.x lmin.y lmin.z lmax.x
//   n1 = lmax.y lmax.z rmin.x rmin.y
//   n2 = rmin.z rmax.x rmax.y rmax.z
//   n3 = left-ref right-ref parent -      (bit patterns, see layout.h)
// The first n_top nodes (breadth-first) are staged in shared memory by the
// kernels; `top` points at that copy (nullptr → everything from global memory).
#pragma once
#include "intersect.h"

namespace rtw {

// deepest Karras hierarchy: 63 key bits + 25 index bits = 88 levels; the 4-wide walk visits every other level and
// pushes at most 3 references per visit: 3 * 44 = 132
#define RTW_STACK 136

struct TravStats {
    uint32_t nodes;
    uint32_t prims;
};

// A push that finds the stack full is COUNTED (DScene::overflow) and dropped; the host reads the word after every
// render / trace call and fails the call (RTW_ERR_STATE) if it is not zero — a wrong image is never returned silently.
// With 63-bit Morton keys and index tie-breaks below 2^25 primitives a Karras hierarchy is at most 63 + 25 = 88
// levels deep, so RTW_STACK = 136 cannot overflow (see the definition) on a hierarchy this library built (tests/test_gpu_edges.py builds
// 4096 coincident primitives to show it).
RTW_HD void push_ref(const DScene& sc, uint32_t* stack, int& sp, uint32_t ref) {
    if (sp < RTW_STACK) {
        stack[sp++] = ref;
    } else if (sc.overflow) {
#if defined(__CUDA_ARCH__)
        atomicAdd(sc.overflow, 1u);
#else
        (*sc.overflow)++;
#endif
    }
}

template <class Mem>
RTW_HD void fetch_node(const DScene& sc, const float4* top, uint32_t ref, float4& n0, float4& n1, float4& n2, float4& n3) {
    if (ref & RTW_REF_TOP) {
        const float4* p = top + 4 * (ref & RTW_REF_MASK);
        n0 = p[0]; n1 = p[1]; n2 = p[2]; n3 = p[3];
    } else {
        const float4* p = sc.nodes + 4 * (size_t)ref;
        n0 = Mem::ld4(p); n1 = Mem::ld4(p + 1); n2 = Mem::ld4(p + 2); n3 = Mem::ld4(p + 3);
    }
}

// Conservative slab test of both children against [tmin, limit].
// t = fma(bound, 1/d, -o/d); any NaN drops out of fminf/fmaxf, which only widens
// the interval.  The far side is padded by 2 ulp-ish (Ize 2013).
struct SlabRay {
    float3 inv, ood;
};
// |d| < 1e-30 is replaced by +-1e-30: with a true 1/0 the fma form gives inf - inf.
// The slab then spans (-huge, +huge) when the origin is inside it, as it should.
RTW_HD float safe_inv(float d) { return 1.0f / (fabsf(d) < 1e-30f ? copysignf(1e-30f, d) : d); }
RTW_HD SlabRay make_slab_ray(const RayW& r) {
    SlabRay s;
    s.inv = f3(safe_inv(r.d.x), safe_inv(r.d.y), safe_inv(r.d.z));
    s.ood = f3(r.o.x * s.inv.x, r.o.y * s.inv.y, r.o.z * s.inv.z);
    return s;
}
RTW_HD bool slab_test(const SlabRay& s, float minx, float miny, float minz, float maxx, float maxy, float maxz,
                      float tmin, float limit, float& tnear) {
    float x0 = fmaf(minx, s.inv.x, -s.ood.x), x1 = fmaf(maxx, s.inv.x, -s.ood.x);
    float y0 = fmaf(miny, s.inv.y, -s.ood.y), y1 = fmaf(maxy, s.inv.y, -s.ood.y);
    float z0 = fmaf(minz, s.inv.z, -s.ood.z), z1 = fmaf(maxz, s.inv.z, -s.ood.z);
    float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tmin));
    float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), limit));
    tnear = tn;
    return tn <= tf * 1.0000004f;
}

// ---- 4-wide nodes (lbvh.h make_wide_node): one visit = one 128-byte record, four slab tests ----
RTW_HD void sort2(float& ta, uint32_t& ra, float& tb, uint32_t& rb) {   // ascending by t
    const bool sw = tb < ta;
    const float t0 = sw ? tb : ta, t1 = sw ? ta : tb;
    const uint32_t r0 = sw ? rb : ra, r1 = sw ? ra : rb;
    ta = t0; tb = t1; ra = r0; rb = r1;
}
// Visits wide node `cur`: returns the child to continue with (the nearest one for a closest-hit walk) after pushing
// the other children the ray enters, nearest on top; RTW_REF_NONE if the ray enters none (the caller pops).
template <class Mem, bool kAny>
RTW_HD uint32_t visit_wide(const DScene& sc, const SlabRay& sr, uint32_t cur, float tmin, float limit, uint32_t* stack, int& sp) {
    const float4* p = sc.wide + 8 * (size_t)cur;
    const float4 mnx = Mem::ld4(p), mny = Mem::ld4(p + 1), mnz = Mem::ld4(p + 2);
    const float4 mxx = Mem::ld4(p + 3), mxy = Mem::ld4(p + 4), mxz = Mem::ld4(p + 5), rf = Mem::ld4(p + 6);
    const float kMiss = 3.0e38f;
    float t0, t1, t2, t3;
    const bool h0 = slab_test(sr, mnx.x, mny.x, mnz.x, mxx.x, mxy.x, mxz.x, tmin, limit, t0);
    const bool h1 = slab_test(sr, mnx.y, mny.y, mnz.y, mxx.y, mxy.y, mxz.y, tmin, limit, t1);
    const bool h2 = slab_test(sr, mnx.z, mny.z, mnz.z, mxx.z, mxy.z, mxz.z, tmin, limit, t2);
    const bool h3 = slab_test(sr, mnx.w, mny.w, mnz.w, mxx.w, mxy.w, mxz.w, tmin, limit, t3);
    uint32_t r0 = f2u(rf.x), r1 = f2u(rf.y), r2 = f2u(rf.z), r3 = f2u(rf.w);
    // a child the ray misses (or that does not exist) sorts behind every child it enters
    t0 = (h0 && r0 != RTW_REF_NONE) ? t0 : kMiss;
    t1 = (h1 && r1 != RTW_REF_NONE) ? t1 : kMiss;
    t2 = (h2 && r2 != RTW_REF_NONE) ? t2 : kMiss;
    t3 = (h3 && r3 != RTW_REF_NONE) ? t3 : kMiss;
    const int hits = (t0 < kMiss ? 1 : 0) + (t1 < kMiss ? 1 : 0) + (t2 < kMiss ? 1 : 0) + (t3 < kMiss ? 1 : 0);
    if (hits == 0) return RTW_REF_NONE;
    // one child entered (the common case near the leaves): nothing to order, nothing to push.  Measured with
    // tools/gpu_ab_variant.sh: +3.9 % on the Book-1 scene (streaming wavefront), +0.8 % on scene 4 and the stress scene.
    if (hits == 1) return t0 < kMiss ? r0 : (t1 < kMiss ? r1 : (t2 < kMiss ? r2 : r3));
    // (a second shortcut for two entered children — one comparison instead of the network — measured -2.8 % on the Book-1
    // scene and +0.5..1 % elsewhere: mixed warps then run three code paths; not kept)
    // (an any-hit walk needs no order, but the sort also compacts the entered children to the front)
    sort2(t0, r0, t1, r1); sort2(t2, r2, t3, r3); sort2(t0, r0, t2, r2); sort2(t1, r1, t3, r3); sort2(t1, r1, t2, r2);
    if (hits > 3) push_ref(sc, stack, sp, r3);
    if (hits > 2) push_ref(sc, stack, sp, r2);
    if (hits > 1) push_ref(sc, stack, sp, r1);
    return r0;
}

RTW_HD int leaf_prim_of(const DScene& sc, uint32_t leaf) {
#if defined(__CUDA_ARCH__)
    return __ldg(sc.leaf_prim + leaf);
#else
    return sc.leaf_prim[leaf];
#endif
}

// Both traversals are "while-while" loops (Aila & Laine 2009): an inner loop over
// internal nodes until the thread holds a leaf range, then an inner loop over the
// primitives of that range.  Lanes of a warp therefore spend their time in the same
// kind of work, and a scene that is a single range (<= leaf_size primitives, e.g. the
// Cornell box) is walked by all 32 lanes in lockstep.

// Closest hit: lexicographic minimum of (t, primitive index).
template <class Mem, bool kStats>
RTW_HD bool traverse_closest(const DScene& sc, const float4* top, const RayW& r, float& t_best, uint32_t& leaf_best,
                             TravStats* stats) {
    SlabRay sr;
    if (!(sc.root & RTW_REF_LEAF)) sr = make_slab_ray(r);   // a single-range scene never tests a box
    XfRay xfc;
    xfc.xf = -1;
    uint32_t stack[RTW_STACK];
    int sp = 0;
    uint32_t cur = sc.root;
    bool found = false;
    float tb = r.tmax;
    uint32_t lb = RTW_REF_NONE;
    // the oversize primitives (lbvh.h) are outside the hierarchy: tested first, by every ray — a warp-uniform loop,
    // and an early limit tb for the walk
    for (uint32_t leaf = (uint32_t)sc.over_first, end = leaf + (uint32_t)sc.over_count; leaf < end; leaf++) {
        float t;
        if (kStats) stats->prims++;
        if (prim_candidate(SrcScene<Mem>{sc}, leaf, r, xfc, t)) {
            if (!found || t < tb) {
                found = true;
                tb = t;
                lb = leaf;
            } else if (t == tb && leaf_prim_of(sc, leaf) < leaf_prim_of(sc, lb)) {
                lb = leaf;
            }
        }
    }
    for (;;) {
        bool exhausted = false;
        while (!(cur & RTW_REF_LEAF)) {
            if (sc.wide) {
                if (kStats) stats->nodes++;
                cur = visit_wide<Mem, false>(sc, sr, cur, r.tmin, tb, stack, sp);
                if (cur == RTW_REF_NONE) {
                    if (sp == 0) { exhausted = true; break; }
                    cur = stack[--sp];
                }
                continue;
            }
            float4 n0, n1, n2, n3;
            fetch_node<Mem>(sc, top, cur, n0, n1, n2, n3);
            if (kStats) stats->nodes++;
            float tl, tr;
            const bool hl = slab_test(sr, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, r.tmin, tb, tl);
            const bool hr = slab_test(sr, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, r.tmin, tb, tr);
            const uint32_t cl = f2u(n3.x), cr = f2u(n3.y);
            if (hl && hr) {
                const bool left_first = tl <= tr;
                push_ref(sc, stack, sp, left_first ? cr : cl);
                cur = left_first ? cl : cr;
            } else if (hl) {
                cur = cl;
            } else if (hr) {
                cur = cr;
            } else {
                if (sp == 0) { exhausted = true; break; }
                cur = stack[--sp];
            }
        }
        if (exhausted) break;
        const uint32_t first = leaf_range_first(cur), end = first + leaf_range_count(cur);
        for (uint32_t leaf = first; leaf < end; leaf++) {
            float t;
            if (kStats) stats->prims++;
            if (prim_candidate(SrcScene<Mem>{sc}, leaf, r, xfc, t)) {
                if (!found || t < tb) {
                    found = true;
                    tb = t;
                    lb = leaf;
                } else if (t == tb && leaf_prim_of(sc, leaf) < leaf_prim_of(sc, lb)) {   // rare tie: lower primitive id wins
                    lb = leaf;
                }
            }
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
    t_best = tb;
    leaf_best = lb;
    return found;
}

// Any hit (shadow rays): true iff some primitive reports an accepted hit in [tmin, tmax].
template <class Mem, bool kStats>
RTW_HD bool traverse_any(const DScene& sc, const float4* top, const RayW& r, TravStats* stats) {
    SlabRay sr;
    if (!(sc.root & RTW_REF_LEAF)) sr = make_slab_ray(r);
    XfRay xfc;
    xfc.xf = -1;
    uint32_t stack[RTW_STACK];
    int sp = 0;
    uint32_t cur = sc.root;
    for (uint32_t leaf = (uint32_t)sc.over_first, end = leaf + (uint32_t)sc.over_count; leaf < end; leaf++) {   // oversize list first
        float t;
        if (kStats) stats->prims++;
        if (prim_candidate(SrcScene<Mem>{sc}, leaf, r, xfc, t)) return true;
    }
    for (;;) {
        bool exhausted = false;
        while (!(cur & RTW_REF_LEAF)) {
            if (sc.wide) {
                if (kStats) stats->nodes++;
                cur = visit_wide<Mem, true>(sc, sr, cur, r.tmin, r.tmax, stack, sp);
                if (cur == RTW_REF_NONE) {
                    if (sp == 0) { exhausted = true; break; }
                    cur = stack[--sp];
                }
                continue;
            }
            float4 n0, n1, n2, n3;
            fetch_node<Mem>(sc, top, cur, n0, n1, n2, n3);
            if (kStats) stats->nodes++;
            float tl, tr;
            const bool hl = slab_test(sr, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, r.tmin, r.tmax, tl);
            const bool hr = slab_test(sr, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, r.tmin, r.tmax, tr);
            const uint32_t cl = f2u(n3.x), cr = f2u(n3.y);
            if (hl && hr) {
                push_ref(sc, stack, sp, cr);
                cur = cl;
            } else if (hl) {
                cur = cl;
            } else if (hr) {
                cur = cr;
            } else {
                if (sp == 0) { exhausted = true; break; }
                cur = stack[--sp];
            }
        }
        if (exhausted) break;
        const uint32_t first = leaf_range_first(cur), end = first + leaf_range_count(cur);
        for (uint32_t leaf = first; leaf < end; leaf++) {
            float t;
            if (kStats) stats->prims++;
            if (prim_candidate(SrcScene<Mem>{sc}, leaf, r, xfc, t)) return true;
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
    return false;
}

// Small scenes (layout.h SmallScene): the one leaf range as a static table plus runs.  Every counter
// here is warp-uniform (it comes from the kernel-parameter copy of the scene): a group / run is "carry
// the ray into its transform once, then one candidate test per record with no dispatch inside", and the
// closest-hit update is selects only.  Same candidates, same lexicographic (t, primitive id) minimum as
// traverse_closest; the order of the tests is free.
struct SmallBest {
    float tb;        // best t so far (r.tmax until something is found)
    uint32_t leaf;   // scene leaf index of the best candidate, RTW_REF_NONE = none
    int32_t pid;     // its primitive id (INT32_MAX until something is found: decides exact ties)
};
template <bool kAny>
RTW_HD void small_accept(bool ok, float t, uint32_t leaf, int32_t pid, SmallBest& best, bool& any) {
    if (kAny) {
        any |= ok;
    } else {
        // a candidate has t <= r.tmax; with nothing found yet tb == r.tmax and pid == INT32_MAX, so t == tmax is taken
        const bool take = ok & ((t < best.tb) | ((t == best.tb) & (pid < best.pid)));
        best.tb = take ? t : best.tb;
        best.leaf = take ? leaf : best.leaf;
        best.pid = take ? pid : best.pid;
    }
}
// The static table, fully unrolled: g, axis and slot are compile-time constants.
template <bool kAny, int kGroup>
RTW_HD void small_group(const SmallScene& ss, const RayW& r, SmallBest& best, bool& any) {
    const SmallGroup& g = ss.group[kGroup];
    float3 o = r.o, d = r.d;
    if (kGroup > 0) {
        o = xform_point(g.w2o[0], g.w2o[1], g.w2o[2], r.o);
        d = xform_vector(g.w2o[0], g.w2o[1], g.w2o[2], r.d);
    }
#pragma unroll
    for (int j = 0; j < RTW_SS_RECTS; j++) {
        if (j >= g.n_rect[0]) break;
        float t;
        const bool ok = rect_candidate(g.rect[0][j].a, g.rect[0][j].k, o.x, d.x, o.y, d.y, o.z, d.z, r.tmin, r.tmax, t);
        small_accept<kAny>(ok, t, (uint32_t)g.rect[0][j].leaf, g.rect[0][j].prim_id, best, any);
    }
#pragma unroll
    for (int j = 0; j < RTW_SS_RECTS; j++) {
        if (j >= g.n_rect[1]) break;
        float t;
        const bool ok = rect_candidate(g.rect[1][j].a, g.rect[1][j].k, o.y, d.y, o.x, d.x, o.z, d.z, r.tmin, r.tmax, t);
        small_accept<kAny>(ok, t, (uint32_t)g.rect[1][j].leaf, g.rect[1][j].prim_id, best, any);
    }
#pragma unroll
    for (int j = 0; j < RTW_SS_RECTS; j++) {
        if (j >= g.n_rect[2]) break;
        float t;
        const bool ok = rect_candidate(g.rect[2][j].a, g.rect[2][j].k, o.z, d.z, o.x, d.x, o.y, d.y, r.tmin, r.tmax, t);
        small_accept<kAny>(ok, t, (uint32_t)g.rect[2][j].leaf, g.rect[2][j].prim_id, best, any);
    }
#pragma unroll
    for (int j = 0; j < RTW_SS_SPHERES; j++) {
        if (j >= g.n_sphere) break;
        float t = 0.f;
        const bool ok = sphere_candidate(f3(g.sphere[j].c), g.sphere[j].c.w, o, d, r.tmin, r.tmax, t);
        small_accept<kAny>(ok, t, (uint32_t)g.sphere[j].leaf, g.sphere[j].prim_id, best, any);
    }
}
// One run of the records that are not in the static table.
template <bool kAny>
RTW_HD void small_run(const SmallScene& ss, const SmallRun& run, const RayW& r, SmallBest& best, bool& any) {
    const SrcSmall src{ss};
    const uint32_t first = (uint32_t)run.first, end = (uint32_t)(run.first + run.count);
    const int kind = run.kind;
    float3 o = r.o, d = r.d;
    if (run.xf >= 0) {
        float4 r0, r1, r2;
        src.w2o(run.xf, r0, r1, r2);
        o = xform_point(r0, r1, r2, r.o);
        d = xform_vector(r0, r1, r2, r.d);
    }
    // (unroll 1: an unrolled body makes the compiler index the constant bank per lane instead of
    // through the uniform datapath)
#pragma unroll 1
    for (uint32_t i = first; i < end; i++) {
        const float4 a = ss.leaf[i].a, b = ss.leaf[i].b;
        float t = 0.f;
        bool ok;
        if (kind == RTW_PRIM_RECT_X) ok = rect_candidate(a, b.x, o.x, d.x, o.y, d.y, o.z, d.z, r.tmin, r.tmax, t);
        else if (kind == RTW_PRIM_RECT_Y) ok = rect_candidate(a, b.x, o.y, d.y, o.x, d.x, o.z, d.z, r.tmin, r.tmax, t);
        else if (kind == RTW_PRIM_RECT_Z) ok = rect_candidate(a, b.x, o.z, d.z, o.x, d.x, o.y, d.y, r.tmin, r.tmax, t);
        else if (kind == RTW_PRIM_SPHERE) ok = sphere_candidate(f3(a), a.w, o, d, r.tmin, r.tmax, t);
        else if (kind == RTW_PRIM_MOVING_SPHERE)
            ok = sphere_candidate(moving_center(a, b, ss.leaf[i].ext, r.gather), a.w, r.o - motion_translation(a, b, r.time), r.d,
                                  r.tmin, r.tmax, t);
        else ok = medium_candidate(kind == RTW_PRIM_VOLUME_BOX, a, b, ss.leaf[i].prim_id, o, d, r, t);
        small_accept<kAny>(ok, t, (uint32_t)ss.leaf[i].leaf, ss.leaf[i].prim_id, best, any);
    }
}
template <bool kAny>
RTW_HD bool small_walk(const SmallScene& ss, const RayW& r, SmallBest& best) {
    best.tb = r.tmax;
    best.leaf = RTW_REF_NONE;
    best.pid = 0x7fffffff;
    bool any = false;
#if defined(__CUDA_ARCH__)
#define RTW_SMALL_ALL_DONE() (kAny && __all_sync(0xffffffffu, any))
#else
#define RTW_SMALL_ALL_DONE() (kAny && any)
#endif
    // any hit: a lane that has its answer keeps walking (its result does not change any more) until
    // every lane of the warp has one; the vote is taken once per group / run
    if (0 < ss.n_groups) small_group<kAny, 0>(ss, r, best, any);
    if (1 < ss.n_groups && !RTW_SMALL_ALL_DONE()) small_group<kAny, 1>(ss, r, best, any);
    if (2 < ss.n_groups && !RTW_SMALL_ALL_DONE()) small_group<kAny, 2>(ss, r, best, any);
    for (int ri = 0; ri < ss.n_runs; ri++) {
        if (RTW_SMALL_ALL_DONE()) break;
        small_run<kAny>(ss, ss.run[ri], r, best, any);
    }
#undef RTW_SMALL_ALL_DONE
    return kAny ? any : best.leaf != RTW_REF_NONE;
}
// leaf_best: index in the scene's leaf order, as traverse_closest reports it
RTW_HD bool small_closest(const SmallScene& ss, const RayW& r, float& t_best, uint32_t& leaf_best, int32_t* prim_best = nullptr) {
    SmallBest best;
    const bool found = small_walk<false>(ss, r, best);
    t_best = best.tb;
    leaf_best = best.leaf;
    if (prim_best) *prim_best = found ? best.pid : -1;
    return found;
}
RTW_HD bool small_any(const SmallScene& ss, const RayW& r) {
    SmallBest best;
    return small_walk<true>(ss, r, best);
}

}  // namespace rtw

// ---------------------------------------------------------------------------
// Resumable traversal ("walker"): the same walk as traverse_closest / traverse_any, cut
// into steps so that a persistent warp can (i) bound how long its lanes run apart — at
// most max_nodes node visits, then at most max_prims primitive tests per step — and
// (ii) hand a finished lane a new ray while the others are still walking (dynamic ray
// fetching, Aila & Laine 2009).  The result does not depend on the step sizes: closest
// hit is the lexicographic minimum of (t, primitive id), any hit is an existence test.
namespace rtw {

template <class Mem, bool kAny>
struct Walker {
    RayW r;
    SlabRay sr;
    XfRay xfc;
    uint32_t cur;        // reference to process next (valid when no range is in progress)
    uint32_t li, lend;   // leaf range in progress: [li, lend); li == lend → none
    int sp;
    float tb;            // closest: best t so far (culling limit); any: r.tmax
    uint32_t lb;
    bool found, done;
    // the stack lives OUTSIDE the struct (a dynamically indexed member would force every field
    // of the walker into local memory); the caller owns `uint32_t stack[RTW_STACK]`

    RTW_HD void begin(const DScene& sc, const RayW& ray, uint32_t* stack) {
        r = ray;
        if (!(sc.root & RTW_REF_LEAF)) sr = make_slab_ray(r);
        xfc.xf = -1;
        cur = sc.root;
        li = lend = 0;
        sp = 0;
        if (sc.over_count > 0) {   // the oversize list (lbvh.h) is the first range in progress; the root waits on the stack
            li = (uint32_t)sc.over_first;
            lend = li + (uint32_t)sc.over_count;
            stack[sp++] = sc.root;
        }
        tb = r.tmax;
        lb = RTW_REF_NONE;
        found = false;
        done = false;
    }

    RTW_HD void pop(const uint32_t* stack) {
        if (sp == 0) done = true;
        else cur = stack[--sp];
    }

    RTW_HD void visit_node(const DScene& sc, const float4* top, uint32_t* stack) {
        if (sc.wide) {
            const uint32_t next = visit_wide<Mem, kAny>(sc, sr, cur, r.tmin, tb, stack, sp);
            if (next == RTW_REF_NONE) pop(stack);
            else cur = next;
            return;
        }
        float4 n0, n1, n2, n3;
        fetch_node<Mem>(sc, top, cur, n0, n1, n2, n3);
        float tl, tr;
        const bool hl = slab_test(sr, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, r.tmin, tb, tl);
        const bool hr = slab_test(sr, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, r.tmin, tb, tr);
        const uint32_t cl = f2u(n3.x), cr = f2u(n3.y);
        if (hl && hr) {
            const bool left_first = kAny || tl <= tr;
            push_ref(sc, stack, sp, left_first ? cr : cl);
            cur = left_first ? cl : cr;
        } else if (hl) {
            cur = cl;
        } else if (hr) {
            cur = cr;
        } else {
            pop(stack);
        }
    }
    RTW_HD void test_prim(const DScene& sc, const uint32_t* stack) {
        float t;
        if (prim_candidate(SrcScene<Mem>{sc}, li, r, xfc, t)) {
            if (kAny) {
                found = true;
                done = true;
            } else if (!found || t < tb) {
                found = true;
                tb = t;
                lb = li;
            } else if (t == tb && leaf_prim_of(sc, li) < leaf_prim_of(sc, lb)) {
                lb = li;
            }
        }
        li++;
        if (!done && li == lend) pop(stack);   // range consumed
    }
    RTW_HD bool wants_node() const { return !done && li == lend && !(cur & RTW_REF_LEAF); }
    RTW_HD void open_leaf() {
        if (!done && li == lend && (cur & RTW_REF_LEAF)) {
            li = leaf_range_first(cur);
            lend = li + leaf_range_count(cur);
        }
    }

    RTW_HD void step(const DScene& sc, const float4* top, uint32_t* stack, int max_nodes, int max_prims) {
        int n = 0;
        while (wants_node() && n < max_nodes) {
            visit_node(sc, top, stack);
            n++;
        }
        open_leaf();
        int p = 0;
        while (!done && li < lend && p < max_prims) {
            test_prim(sc, stack);
            p++;
        }
    }

#if defined(__CUDACC__)
    // The same step, called by ALL 32 lanes of a converged warp (a lane without a ray has done == true), with
    // the two inner loops under a warp vote: the node loop stops as soon as fewer than `node_min` lanes still
    // have an internal node to visit (the others have reached a leaf and wait), the primitive loop as soon as
    // fewer than `prim_min` lanes still have a primitive to test.  Lanes that are cut short carry on in the next
    // step; the result of a walk does not depend on where it is cut.  (ncu, scene 4, plain step: 5.4 of 32 lanes
    // active in the slab tests — a few long descents kept the loop going for everyone.)
    __device__ __forceinline__ void step_voted(const DScene& sc, const float4* top, uint32_t* stack, int max_nodes, int max_prims,
                                               int node_min, int prim_min) {
        for (int n = 0; n < max_nodes; n++) {
            const bool want = wants_node();
            const int k = __popc(__ballot_sync(0xffffffffu, want));
            if (k == 0 || (n > 0 && k < node_min)) break;
            if (want) visit_node(sc, top, stack);
        }
        open_leaf();
        for (int p = 0; p < max_prims; p++) {
            const bool want = !done && li < lend;
            const int k = __popc(__ballot_sync(0xffffffffu, want));
            if (k == 0 || (p > 0 && k < prim_min)) break;
            if (want) test_prim(sc, stack);
        }
    }
#endif
};

}  // namespace rtw
