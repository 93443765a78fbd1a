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

#define RTW_STACK 96

struct TravStats {
    uint32_t nodes;
    uint32_t prims;
};

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
    for (;;) {
        bool exhausted = false;
        while (!(cur & RTW_REF_LEAF)) {
            float4 n0, n1, n2, n3;
            fetch_node<Mem>(sc, top, cur, n0, n1, n2, n3);
            if (kStats) stats->nodes++;
            float tl, tr;
            const bool hl = slab_test(sr, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, r.tmin, tb, tl);
            const bool hr = slab_test(sr, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, r.tmin, tb, tr);
            const uint32_t cl = f2u(n3.x), cr = f2u(n3.y);
            if (hl && hr) {
                const bool left_first = tl <= tr;
                if (sp < RTW_STACK) stack[sp++] = left_first ? cr : cl;
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
    for (;;) {
        bool exhausted = false;
        while (!(cur & RTW_REF_LEAF)) {
            float4 n0, n1, n2, n3;
            fetch_node<Mem>(sc, top, cur, n0, n1, n2, n3);
            if (kStats) stats->nodes++;
            float tl, tr;
            const bool hl = slab_test(sr, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, r.tmin, r.tmax, tl);
            const bool hr = slab_test(sr, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, r.tmin, r.tmax, tr);
            const uint32_t cl = f2u(n3.x), cr = f2u(n3.y);
            if (hl && hr) {
                if (sp < RTW_STACK) stack[sp++] = cr;
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

// Small scenes (layout.h SmallScene): the one leaf range, leaf by leaf, with a warp-uniform
// counter.  Same candidates, same lexicographic (t, primitive id) minimum as traverse_closest.
RTW_HD bool small_closest(const SmallScene& ss, const RayW& r, float& t_best, uint32_t& leaf_best) {
    const SrcSmall src{ss};
    XfRay cur;
    cur.xf = -1;
    cur.o = r.o;
    cur.d = r.d;
    bool found = false;
    float tb = r.tmax;
    uint32_t lb = RTW_REF_NONE;
    for (int leaf = 0; leaf < ss.n; leaf++) {
        float t;
        if (prim_candidate_uniform(src, (uint32_t)leaf, r, cur, t)) {
            if (!found || t < tb) {
                found = true;
                tb = t;
                lb = (uint32_t)leaf;
            } else if (t == tb && ss.leaf[leaf].prim_id < ss.leaf[lb].prim_id) {
                lb = (uint32_t)leaf;
            }
        }
    }
    t_best = tb;
    leaf_best = lb;
    return found;
}
// Any hit.  The loop stays warp-uniform: a lane that has its answer keeps walking (its result is
// not changed any more) until every lane of the warp has one.
RTW_HD bool small_any(const SmallScene& ss, const RayW& r) {
    const SrcSmall src{ss};
    XfRay cur;
    cur.xf = -1;
    cur.o = r.o;
    cur.d = r.d;
    bool hit = false;
    for (int leaf = 0; leaf < ss.n; leaf++) {
        float t;
        if (prim_candidate_uniform(src, (uint32_t)leaf, r, cur, t)) hit = true;
#if defined(__CUDA_ARCH__)
        if (__all_sync(0xffffffffu, hit)) break;
#else
        if (hit) break;
#endif
    }
    return hit;
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

    RTW_HD void begin(const DScene& sc, const RayW& ray) {
        r = ray;
        if (!(sc.root & RTW_REF_LEAF)) sr = make_slab_ray(r);
        xfc.xf = -1;
        cur = sc.root;
        li = lend = 0;
        sp = 0;
        tb = r.tmax;
        lb = RTW_REF_NONE;
        found = false;
        done = false;
    }

    RTW_HD void pop(const uint32_t* stack) {
        if (sp == 0) done = true;
        else cur = stack[--sp];
    }

    RTW_HD void step(const DScene& sc, const float4* top, uint32_t* stack, int max_nodes, int max_prims) {
        int n = 0;
        while (!done && li == lend && !(cur & RTW_REF_LEAF) && n < max_nodes) {
            float4 n0, n1, n2, n3;
            fetch_node<Mem>(sc, top, cur, n0, n1, n2, n3);
            float tl, tr;
            const bool hl = slab_test(sr, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, r.tmin, tb, tl);
            const bool hr = slab_test(sr, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, r.tmin, tb, tr);
            const uint32_t cl = f2u(n3.x), cr = f2u(n3.y);
            if (hl && hr) {
                const bool left_first = kAny || tl <= tr;
                if (sp < RTW_STACK) stack[sp++] = left_first ? cr : cl;
                cur = left_first ? cl : cr;
            } else if (hl) {
                cur = cl;
            } else if (hr) {
                cur = cr;
            } else {
                pop(stack);
            }
            n++;
        }
        if (!done && li == lend && (cur & RTW_REF_LEAF)) {
            li = leaf_range_first(cur);
            lend = li + leaf_range_count(cur);
        }
        int p = 0;
        while (!done && li < lend && p < max_prims) {
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
            p++;
            if (!done && li == lend) pop(stack);   // range consumed
        }
    }
};

}  // namespace rtw
