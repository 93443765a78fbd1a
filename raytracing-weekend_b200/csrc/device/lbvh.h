// lbvh.h — per-thread bodies of the GPU LBVH builder.
//
// Replaces every optixAccelBuild call site of the reference (one single-primitive
// GAS per primitive plus one IAS: geometry/ioSphere.h:45-92, ioAARect.h:96,
// ioMovingSphere.h:135, ioVolumeBox.h:69, ioVolumeSphere.h:69,
// ioGeometryGroup.h:160-225) by ONE two-child LBVH over all primitives:
//   world AABB per primitive → scene bounds → 63-bit Morton code of the centroid
//   → stable radix sort → Karras 2012 hierarchy → bottom-up refit.
// Primitive AABBs follow the reference's bounds functions (cited below) and are
// padded by 2^-21 of the largest |coordinate| so that the slab test can never
// cull a primitive whose intersection test would accept the ray.
#pragma once
#include "layout.h"

namespace rtw {

struct Box3 {
    float3 mn, mx;
};

RTW_HD Box3 prim_object_aabb(const rtw_prim& p) {
    Box3 b;
    switch (p.type) {
    case RTW_PRIM_SPHERE: {  // geometry/ioSphere.h:25-34
        float3 c = f3(p.v[0], p.v[1], p.v[2]);
        b.mn = c - p.v[3];
        b.mx = c + p.v[3];
        break;
    }
    case RTW_PRIM_MOVING_SPHERE: {  // geometry/ioMovingSphere.h:46-72, swept by the motion keys (Q24)
        float3 c0 = f3(p.v[0], p.v[1], p.v[2]), c1 = f3(p.v[4], p.v[5], p.v[6]);
        float r = p.v[3];
        float3 lo = vmin(c0 - r, c1 - r), hi = vmax(c0 + r, c1 + r);
        b.mn = vmin(lo + c0, lo + c1);
        b.mx = vmax(hi + c0, hi + c1);
        break;
    }
    case RTW_PRIM_RECT_X:  // geometry/ioAARect.h:128-154
        b.mn = f3(p.v[4] - 0.0001f, p.v[0], p.v[2]);
        b.mx = f3(p.v[4] + 0.0001f, p.v[1], p.v[3]);
        break;
    case RTW_PRIM_RECT_Y:
        b.mn = f3(p.v[0], p.v[4] - 0.0001f, p.v[2]);
        b.mx = f3(p.v[1], p.v[4] + 0.0001f, p.v[3]);
        break;
    case RTW_PRIM_RECT_Z:
        b.mn = f3(p.v[0], p.v[2], p.v[4] - 0.0001f);
        b.mx = f3(p.v[1], p.v[3], p.v[4] + 0.0001f);
        break;
    case RTW_PRIM_VOLUME_BOX:  // geometry/ioVolumeBox.h:76-84
        b.mn = f3(p.v[0], p.v[1], p.v[2]) - 0.0001f;
        b.mx = f3(p.v[3], p.v[4], p.v[5]) + 0.0001f;
        break;
    default: {  // geometry/ioVolumeSphere.h:76-82
        float3 c = f3(p.v[0], p.v[1], p.v[2]);
        float r = fabsf(p.v[3]);
        b.mn = c - r;
        b.mx = c + r;
        break;
    }
    }
    return b;
}

// World-space, padded AABB of primitive p (instance transform applied to the 8 corners).
RTW_HD Box3 prim_world_aabb(const rtw_prim& p, const rtw_transform* xforms) {
    Box3 b = prim_object_aabb(p);
    if (p.type != RTW_PRIM_MOVING_SPHERE && p.transform >= 0) {
        const float* m = xforms[p.transform].o2w;
        const float4 r0 = make_float4(m[0], m[1], m[2], m[3]);
        const float4 r1 = make_float4(m[4], m[5], m[6], m[7]);
        const float4 r2 = make_float4(m[8], m[9], m[10], m[11]);
        Box3 w;
        w.mn = f3(RTW_FLT_MAX);
        w.mx = f3(-RTW_FLT_MAX);
        for (int i = 0; i < 8; i++) {
            float3 c = f3((i & 1) ? b.mx.x : b.mn.x, (i & 2) ? b.mx.y : b.mn.y, (i & 4) ? b.mx.z : b.mn.z);
            float3 q = xform_point(r0, r1, r2, c);
            w.mn = vmin(w.mn, q);
            w.mx = vmax(w.mx, q);
        }
        b = w;
    }
    float m = fmaxf(fmaxf(fmaxf(fabsf(b.mn.x), fabsf(b.mn.y)), fabsf(b.mn.z)),
                    fmaxf(fmaxf(fabsf(b.mx.x), fabsf(b.mx.y)), fabsf(b.mx.z)));
    float pad = m * 4.76837158203125e-7f;  // 2^-21
    b.mn = b.mn - pad;
    b.mx = b.mx + pad;
    return b;
}

// ---- oversize primitives ----------------------------------------------------------------------------
// A primitive whose box is large against the spread of the CENTROIDS, and whose own centroid lies among the small
// ones, sits deep in a Morton-ordered hierarchy and inflates every ancestor box up to the root: every ray then walks
// that whole chain.  (Book-2 final scene, scene/ioScene.h:838,931: the r = 5000 shell and the r = 500 fog sphere
// around the origin, amid 2400 box faces of size 100 — 5.7 of the 20.7 node visits per ray, oracle counters.)
// Such primitives (box extent > 1/4 of the largest extent of the centroid bounds; at most RTW_OVER_MAX of them,
// and at least two others left) are kept OUT of the hierarchy: their keys get bit 63, so the sort puts them behind
// everything else, the hierarchy is built over the first tree_leaf_count() leaves only, and every ray tests the
// last leaves [m, n) directly before it walks.  The closest hit is the lexicographic minimum of (t, primitive id):
// where a candidate is tested does not change it.  Flagged keys keep bit 63 also when the list is not used
// (too many flagged, too few others): the hierarchy then simply covers all leaves.
#define RTW_OVER_MAX 8
#define RTW_KEY_OVERSIZE (1ull << 63)
RTW_HD float3 box_centroid(const Box3& b) { return (b.mn + b.mx) * 0.5f; }
RTW_HD bool box_is_oversize(const Box3& b, float3 cen_lo, float3 cen_hi) {
    const float3 ce = cen_hi - cen_lo, be = b.mx - b.mn;
    return fmaxf(fmaxf(be.x, be.y), be.z) > 0.25f * fmaxf(fmaxf(ce.x, ce.y), ce.z);
}
RTW_HD int tree_leaf_count(int n, int flagged) {
    return (flagged > 0 && flagged <= RTW_OVER_MAX && n - flagged >= 2) ? n - flagged : n;
}

RTW_HD unsigned long long expand21(uint32_t v) {
    unsigned long long x = v & 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}
RTW_HD uint32_t quant21(float c, float lo, float ext) {
    float n = (ext > 0.f) ? (c - lo) / ext : 0.f;
    float q = n * 2097152.0f;
    if (!(q >= 0.f)) q = 0.f;
    if (q > 2097151.0f) q = 2097151.0f;
    return (uint32_t)q;
}
RTW_HD unsigned long long morton63(const Box3& b, float3 lo, float3 ext) {
    float3 c = box_centroid(b);
    return (expand21(quant21(c.x, lo.x, ext.x)) << 2) | (expand21(quant21(c.y, lo.y, ext.y)) << 1) |
           expand21(quant21(c.z, lo.z, ext.z));
}

RTW_HD int clz64(unsigned long long x) {
#if defined(__CUDA_ARCH__)
    return __clzll((long long)x);
#else
    return x ? __builtin_clzll(x) : 64;
#endif
}
RTW_HD int clz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __clz((int)x);
#else
    return x ? __builtin_clz(x) : 32;
#endif
}

// common-prefix length of keys i and j, ties broken by index (Karras 2012, §4)
RTW_HD int lbvh_delta(const unsigned long long* k, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const unsigned long long a = k[i], b = k[j];
    if (a == b) return 64 + clz32((uint32_t)i ^ (uint32_t)j);
    return clz64(a ^ b);
}

// Internal node i of the hierarchy over n >= 2 sorted keys.  Child refs use the
// download convention: >= 0 internal node, < 0 → leaf ~ref.  The node spans the
// leaves [lo, hi]; its left child spans [lo, gamma], its right child [gamma+1, hi].
RTW_HD void karras_node(const unsigned long long* k, int n, int i, int& left, int& right, int& lo_out, int& gamma_out,
                        int& hi_out) {
    const int d = (lbvh_delta(k, n, i, i + 1) - lbvh_delta(k, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = lbvh_delta(k, n, i, i - d);
    int lmax = 2;
    while (lbvh_delta(k, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (lbvh_delta(k, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = lbvh_delta(k, n, i, j);
    int sp = 0;
    int t = l;
    do {
        t = (t + 1) >> 1;
        if (lbvh_delta(k, n, i, i + (sp + t) * d) > dnode) sp += t;
    } while (t > 1);
    const int gamma = i + sp * d + (d < 0 ? d : 0);
    const int lo = i < j ? i : j, hi = i < j ? j : i;
    left = (lo == gamma) ? ~gamma : gamma;
    right = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    lo_out = lo;
    gamma_out = gamma;
    hi_out = hi;
}
RTW_HD void karras_node(const unsigned long long* k, int n, int i, int& left, int& right) {
    int lo, gamma, hi;
    karras_node(k, n, i, left, right, lo, gamma, hi);
}

// Convert a download-convention ref into the traversal encoding of layout.h.
RTW_HD uint32_t encode_ref(int ref) { return ref < 0 ? (RTW_REF_LEAF | (uint32_t)(~ref)) : (uint32_t)ref; }
// Traversal ref of the child spanning leaves [first, last]: one leaf range when the
// subtree holds at most leaf_size primitives, else the internal node itself.
RTW_HD uint32_t traversal_ref(int karras_ref, int first, int last, int leaf_size) {
    const int count = last - first + 1;
    if (count <= leaf_size) return leaf_range_ref(first, count);
    return (uint32_t)karras_ref;
}

// ---- 4-wide nodes ---------------------------------------------------------------------------
// Wide node i = two-child node i with every child that is itself an internal node replaced by THAT node's two
// children: 2..4 children, one 128-byte record, half the node visits and stack traffic of the two-child walk.
// It exists for every internal node (a child reference names the two-child node; the walk reads wide[ref]).
//   w[0..5] = min.x min.y min.z max.x max.y max.z of children 0..3 (structure of arrays: one float4 per bound)
//   w[6]    = the 4 child references (layout.h encoding; RTW_REF_NONE = no child, box = unreachable point)
//   w[7]    = unused (keeps the record at 128 bytes)
RTW_HD void node_child(const float4& n0, const float4& n1, const float4& n2, const float4& n3, int side, float3& mn, float3& mx,
                       uint32_t& ref) {
    if (side == 0) { mn = f3(n0.x, n0.y, n0.z); mx = f3(n0.w, n1.x, n1.y); ref = f2u(n3.x); }
    else { mn = f3(n1.z, n1.w, n2.x); mx = f3(n2.y, n2.z, n2.w); ref = f2u(n3.y); }
}
RTW_HD void make_wide_node(const float4* nodes, int i, float4* w) {
    float bx[6][4];
    uint32_t refs[4];
    for (int k = 0; k < 4; k++) {
        for (int c = 0; c < 6; c++) bx[c][k] = RTW_FLT_MAX;
        refs[k] = RTW_REF_NONE;
    }
    int m = 0;
    const float4 a0 = nodes[4 * i], a1 = nodes[4 * i + 1], a2 = nodes[4 * i + 2], a3 = nodes[4 * i + 3];
    for (int side = 0; side < 2; side++) {
        float3 mn, mx;
        uint32_t ref;
        node_child(a0, a1, a2, a3, side, mn, mx, ref);
        if (ref & RTW_REF_LEAF) {
            bx[0][m] = mn.x; bx[1][m] = mn.y; bx[2][m] = mn.z; bx[3][m] = mx.x; bx[4][m] = mx.y; bx[5][m] = mx.z;
            refs[m++] = ref;
        } else {
            const float4 b0 = nodes[4 * ref], b1 = nodes[4 * ref + 1], b2 = nodes[4 * ref + 2], b3 = nodes[4 * ref + 3];
            for (int s2 = 0; s2 < 2; s2++) {
                float3 cmn, cmx;
                uint32_t cref;
                node_child(b0, b1, b2, b3, s2, cmn, cmx, cref);
                bx[0][m] = cmn.x; bx[1][m] = cmn.y; bx[2][m] = cmn.z; bx[3][m] = cmx.x; bx[4][m] = cmx.y; bx[5][m] = cmx.z;
                refs[m++] = cref;
            }
        }
    }
    for (int c = 0; c < 6; c++) w[c] = make_float4(bx[c][0], bx[c][1], bx[c][2], bx[c][3]);
    w[6] = make_float4(u2f(refs[0]), u2f(refs[1]), u2f(refs[2]), u2f(refs[3]));
    w[7] = make_float4(0.f, 0.f, 0.f, 0.f);
}

// Pack primitive p (material type mt) into its two-float4 leaf record.
RTW_HD void pack_prim(const rtw_prim& p, int mattype, float4& a, float4& b, float2& ext) {
    const float meta = u2f(meta_pack(p.type, p.flip, mattype, p.type == RTW_PRIM_MOVING_SPHERE ? -1 : p.transform));
    ext = make_float2(0.f, 0.f);
    switch (p.type) {
    case RTW_PRIM_SPHERE:
        a = make_float4(p.v[0], p.v[1], p.v[2], p.v[3]);
        b = make_float4(0.f, 0.f, 0.f, meta);
        break;
    case RTW_PRIM_MOVING_SPHERE:
        a = make_float4(p.v[0], p.v[1], p.v[2], p.v[3]);
        b = make_float4(p.v[4], p.v[5], p.v[6], meta);
        ext = make_float2(p.v[7], p.v[8]);
        break;
    case RTW_PRIM_RECT_X:
    case RTW_PRIM_RECT_Y:
    case RTW_PRIM_RECT_Z:
        a = make_float4(p.v[0], p.v[1], p.v[2], p.v[3]);
        b = make_float4(p.v[4], 0.f, 0.f, meta);
        break;
    case RTW_PRIM_VOLUME_BOX:
        a = make_float4(p.v[0], p.v[1], p.v[2], p.v[6]);
        b = make_float4(p.v[3], p.v[4], p.v[5], meta);
        break;
    default:
        a = make_float4(p.v[0], p.v[1], p.v[2], p.v[3]);
        b = make_float4(p.v[4], 0.f, 0.f, meta);
        break;
    }
}

}  // namespace rtw
