// TEST HARNESS ONLY — never shipped, never linked into librtw.so.
//
// Host-side driver for the product's per-thread kernel bodies
// (raytracing-weekend_b200/csrc/device/*.h compiled with g++).  It runs the
// SAME functions the CUDA kernels call — LBVH builder bodies, traversal,
// wavefront stage bodies — in plain loops, so that the wavefront state machine,
// the RNG ledger and the LBVH can be checked against the oracle in this
// GPU-less container (`pytest -m "not gpu"`).  What it cannot cover (warp
// ballots, atomics, shared-memory staging, the radix sort kernels) is covered
// by the `-m gpu` tests through the C-ABI.
#include <algorithm>
#include <cstring>
#include <vector>

#include "../../raytracing-weekend_b200/csrc/device/lbvh.h"
#include "../../raytracing-weekend_b200/csrc/device/wavefront.h"

using namespace rtw;

struct Emu {
    std::vector<rtw_prim> prims;
    std::vector<rtw_transform> xforms;
    std::vector<rtw_material> mats_raw;
    // LBVH
    std::vector<unsigned long long> keys;
    std::vector<int> leaf_prim, leaf_mat;
    std::vector<float4> leaf_mn, leaf_mx, packed, nodes, top_nodes, wide;
    std::vector<float2> ext;
    std::vector<int> left, right;
    // tables
    std::vector<float4> xf4, mats, texs, perlin_vec;
    std::vector<int> perlin_perm;
    std::vector<std::vector<uchar4>> image_data;
    std::vector<DImage> images;
    DScene sc{};
    float scene_min[3], scene_max[3];
};

extern "C" Emu* emu_create(const rtw_scene_desc* d, int max_top, int leaf_size) {
    if (leaf_size <= 0) leaf_size = d->n_prims <= 16 ? 16 : 1;   // same default as rtw_build_accel
    if (leaf_size > RTW_MAX_LEAF_SIZE) leaf_size = RTW_MAX_LEAF_SIZE;
    Emu* e = new Emu();
    const int n = d->n_prims;
    e->prims.assign(d->prims, d->prims + n);
    if (d->n_transforms) e->xforms.assign(d->transforms, d->transforms + d->n_transforms);
    e->mats_raw.assign(d->materials, d->materials + d->n_materials);
    DScene& sc = e->sc;
    // ---- tables exactly as rtw_upload_scene lays them out ----
    e->xf4.resize((size_t)6 * std::max(d->n_transforms, 1));
    for (int i = 0; i < d->n_transforms; i++) {
        const rtw_transform& t = d->transforms[i];
        for (int r = 0; r < 3; r++) {
            e->xf4[6 * i + r] = make_float4(t.w2o[4 * r], t.w2o[4 * r + 1], t.w2o[4 * r + 2], t.w2o[4 * r + 3]);
            e->xf4[6 * i + 3 + r] = make_float4(t.o2w[4 * r], t.o2w[4 * r + 1], t.o2w[4 * r + 2], t.o2w[4 * r + 3]);
        }
    }
    sc.xforms = e->xf4.data();
    e->mats.resize(d->n_materials);
    for (int i = 0; i < d->n_materials; i++)
        e->mats[i] = make_float4(u2f((uint32_t)d->materials[i].type), u2f((uint32_t)d->materials[i].texture), d->materials[i].param, 0.f);
    sc.mats = e->mats.data();
    e->texs.resize((size_t)2 * std::max(d->n_textures, 1));
    for (int i = 0; i < d->n_textures; i++) {
        const rtw_texture& t = d->textures[i];
        e->texs[2 * i] = make_float4(u2f((uint32_t)t.type), t.color[0], t.color[1], t.color[2]);
        const int idx = t.type == RTW_TEX_NOISE ? t.perlin : t.image;
        e->texs[2 * i + 1] = make_float4(u2f((uint32_t)t.odd), u2f((uint32_t)t.even), t.scale, u2f((uint32_t)idx));
    }
    sc.texs = e->texs.data();
    e->perlin_vec.resize((size_t)256 * std::max(d->n_perlin, 1));
    e->perlin_perm.resize((size_t)768 * std::max(d->n_perlin, 1));
    for (int t = 0; t < d->n_perlin; t++) {
        const rtw_perlin& p = d->perlin[t];
        for (int i = 0; i < 256; i++) {
            e->perlin_vec[256 * t + i] = make_float4(p.ranvec[i][0], p.ranvec[i][1], p.ranvec[i][2], 0.f);
            e->perlin_perm[768 * t + i] = p.perm_x[i];
            e->perlin_perm[768 * t + 256 + i] = p.perm_y[i];
            e->perlin_perm[768 * t + 512 + i] = p.perm_z[i];
        }
    }
    sc.perlin_vec = e->perlin_vec.data();
    sc.perlin_perm = e->perlin_perm.data();
    for (int i = 0; i < d->n_images; i++) {
        const rtw_image& im = d->images[i];
        std::vector<uchar4> px((size_t)im.width * im.height);
        memcpy(px.data(), im.rgba, px.size() * 4);
        e->image_data.push_back(std::move(px));
    }
    for (int i = 0; i < d->n_images; i++) {
        DImage di;
        di.texels = e->image_data[i].data();
        di.width = d->images[i].width;
        di.height = d->images[i].height;
        e->images.push_back(di);
    }
    sc.images = e->images.empty() ? nullptr : e->images.data();
    sc.has_images = d->n_images > 0;
    for (int i = 0; i < n; i++) {
        if (d->prims[i].type == RTW_PRIM_VOLUME_BOX || d->prims[i].type == RTW_PRIM_VOLUME_SPHERE) sc.has_media = 1;
        if (d->prims[i].type == RTW_PRIM_MOVING_SPHERE) sc.has_motion = 1;
    }
    sc.n_lights = d->n_lights;
    sc.sky_light = d->sky_light;
    sc.has_light_rect = d->has_light_rect;
    sc.rect_axis = d->light_rect.axis;
    sc.rect_a0 = d->light_rect.a0; sc.rect_a1 = d->light_rect.a1;
    sc.rect_b0 = d->light_rect.b0; sc.rect_b1 = d->light_rect.b1;
    sc.rect_k = d->light_rect.k;
    for (int i = 0; i < d->n_lights && i < RTW_MAX_LIGHTS; i++) {
        sc.lights[i].normal = f3(d->lights[i].normal[0], d->lights[i].normal[1], d->lights[i].normal[2]);
        sc.lights[i].area = d->lights[i].area;
        sc.lights[i].emission = f3(d->lights[i].emission[0], d->lights[i].emission[1], d->lights[i].emission[2]);
    }
    const rtw_camera& cm = d->camera;
    sc.cam_origin = f3(cm.origin[0], cm.origin[1], cm.origin[2]);
    sc.cam_u = f3(cm.u[0], cm.u[1], cm.u[2]);
    sc.cam_v = f3(cm.v[0], cm.v[1], cm.v[2]);
    sc.cam_ll = f3(cm.lower_left[0], cm.lower_left[1], cm.lower_left[2]);
    sc.cam_horizontal = f3(cm.horizontal[0], cm.horizontal[1], cm.horizontal[2]);
    sc.cam_vertical = f3(cm.vertical[0], cm.vertical[1], cm.vertical[2]);
    sc.cam_lens_radius = cm.lens_radius;
    sc.cam_time0 = cm.time0;
    sc.cam_time1 = cm.time1;

    // ---- LBVH with the product's builder bodies (k_prim_aabb … k_refit in loop form) ----
    std::vector<Box3> boxes(n);
    float3 lo = f3(RTW_FLT_MAX), hi = f3(-RTW_FLT_MAX);
    for (int i = 0; i < n; i++) {
        boxes[i] = prim_world_aabb(e->prims[i], e->xforms.empty() ? nullptr : e->xforms.data());
        lo = vmin(lo, boxes[i].mn);
        hi = vmax(hi, boxes[i].mx);
    }
    e->scene_min[0] = lo.x; e->scene_min[1] = lo.y; e->scene_min[2] = lo.z;
    e->scene_max[0] = hi.x; e->scene_max[1] = hi.y; e->scene_max[2] = hi.z;
    float3 clo = f3(RTW_FLT_MAX), chi = f3(-RTW_FLT_MAX);   // bounds of the box centroids (k_prim_aabb)
    for (int i = 0; i < n; i++) { clo = vmin(clo, box_centroid(boxes[i])); chi = vmax(chi, box_centroid(boxes[i])); }
    std::vector<std::pair<unsigned long long, int>> kv(n);
    int flagged = 0;
    for (int i = 0; i < n; i++) {   // k_morton: oversize primitives sort behind everything else (lbvh.h)
        unsigned long long key = morton63(boxes[i], lo, hi - lo);
        if (box_is_oversize(boxes[i], clo, chi)) { key |= RTW_KEY_OVERSIZE; flagged++; }
        kv[i] = {key, i};
    }
    const int n_all = n;
    std::stable_sort(kv.begin(), kv.end(), [](const auto& a, const auto& b) { return a.first < b.first; });  // stands in for sort.cu
    e->keys.resize(n);
    e->leaf_prim.resize(n); e->leaf_mat.resize(n);
    e->leaf_mn.resize(n); e->leaf_mx.resize(n);
    e->packed.resize((size_t)2 * n); e->ext.resize(n);
    for (int j = 0; j < n; j++) {
        const int pi = kv[j].second;
        e->keys[j] = kv[j].first;
        const rtw_prim& p = e->prims[pi];
        pack_prim(p, e->mats_raw[p.material].type, e->packed[2 * j], e->packed[2 * j + 1], e->ext[j]);
        e->leaf_prim[j] = pi;
        e->leaf_mat[j] = p.material;
        e->leaf_mn[j] = make_float4(boxes[pi].mn.x, boxes[pi].mn.y, boxes[pi].mn.z, 0.f);
        e->leaf_mx[j] = make_float4(boxes[pi].mx.x, boxes[pi].mx.y, boxes[pi].mx.z, 0.f);
    }
    const int m = tree_leaf_count(n_all, flagged);   // leaves under the hierarchy; [m, n) is the oversize list
    const int ni = m > 1 ? m - 1 : 1;
    e->nodes.assign((size_t)4 * ni, make_float4(0, 0, 0, 0));
    e->left.assign(ni, 0); e->right.assign(ni, 0);
    std::vector<int> parent(ni, -1), leaf_parent(n, -1), flags(ni, 0);
    if (n == 1) {
        const float4 mn = e->leaf_mn[0], mx = e->leaf_mx[0];
        e->nodes[0] = make_float4(mn.x, mn.y, mn.z, mx.x);
        e->nodes[1] = make_float4(mx.y, mx.z, RTW_FLT_MAX, RTW_FLT_MAX);
        e->nodes[2] = make_float4(RTW_FLT_MAX, RTW_FLT_MAX, RTW_FLT_MAX, RTW_FLT_MAX);
        e->nodes[3] = make_float4(u2f(RTW_REF_LEAF), u2f(RTW_REF_LEAF), u2f(RTW_REF_LEAF), u2f(RTW_REF_LEAF));
        e->left[0] = ~0; e->right[0] = ~0;
    } else {
        for (int i = 0; i < ni; i++) {
            int l, r, lo, gamma, hi;
            karras_node(e->keys.data(), m, i, l, r, lo, gamma, hi);
            e->left[i] = l; e->right[i] = r;
            if (l < 0) leaf_parent[~l] = i; else parent[l] = i;
            if (r < 0) leaf_parent[~r] = i; else parent[r] = i;
            e->nodes[4 * i + 3] = make_float4(u2f(traversal_ref(l, lo, gamma, leaf_size)), u2f(traversal_ref(r, gamma + 1, hi, leaf_size)),
                                              u2f(encode_ref(l)), u2f(encode_ref(r)));
        }
        parent[0] = -1;
        auto child_box = [&](uint32_t ref, float3& mn, float3& mx) {
            if (ref & RTW_REF_LEAF) { mn = f3(e->leaf_mn[ref & RTW_REF_MASK]); mx = f3(e->leaf_mx[ref & RTW_REF_MASK]); }
            else {
                const float4 n0 = e->nodes[4 * ref], n1 = e->nodes[4 * ref + 1], n2 = e->nodes[4 * ref + 2];
                mn = vmin(f3(n0.x, n0.y, n0.z), f3(n1.z, n1.w, n2.x));
                mx = vmax(f3(n0.w, n1.x, n1.y), f3(n2.y, n2.z, n2.w));
            }
        };
        for (int leaf = 0; leaf < m; leaf++) {  // k_refit, sequentialised
            int cur = leaf_parent[leaf];
            while (cur >= 0) {
                if (flags[cur]++ == 0) break;
                const float4 n3 = e->nodes[4 * cur + 3];
                float3 lmn, lmx, rmn, rmx;
                child_box(f2u(n3.z), lmn, lmx);
                child_box(f2u(n3.w), rmn, rmx);
                e->nodes[4 * cur + 0] = make_float4(lmn.x, lmn.y, lmn.z, lmx.x);
                e->nodes[4 * cur + 1] = make_float4(lmx.y, lmx.z, rmn.x, rmn.y);
                e->nodes[4 * cur + 2] = make_float4(rmn.z, rmx.x, rmx.y, rmx.z);
                cur = parent[cur];
            }
        }
    }
    // k_build_top in loop form
    const int mt = std::min(max_top, ni);
    std::vector<int> top_index(ni, -1), queue;
    if (mt > 0) { queue.push_back(0); top_index[0] = 0; }
    for (size_t head = 0; head < queue.size(); head++) {
        const float4 n3 = e->nodes[4 * queue[head] + 3];
        const uint32_t c[2] = {f2u(n3.x), f2u(n3.y)};
        for (int k = 0; k < 2; k++)
            if (!(c[k] & RTW_REF_LEAF) && (int)queue.size() < mt && top_index[c[k]] < 0) {
                top_index[c[k]] = (int)queue.size();
                queue.push_back((int)c[k]);
            }
    }
    e->top_nodes.resize((size_t)4 * std::max<size_t>(queue.size(), 1));
    for (size_t j = 0; j < queue.size(); j++) {
        const int g = queue[j];
        float4 n3 = e->nodes[4 * g + 3];
        uint32_t cl = f2u(n3.x), cr = f2u(n3.y);
        if (!(cl & RTW_REF_LEAF) && top_index[cl] >= 0) cl = RTW_REF_TOP | (uint32_t)top_index[cl];
        if (!(cr & RTW_REF_LEAF) && top_index[cr] >= 0) cr = RTW_REF_TOP | (uint32_t)top_index[cr];
        n3.x = u2f(cl); n3.y = u2f(cr);
        for (int k = 0; k < 3; k++) e->top_nodes[4 * j + k] = e->nodes[4 * g + k];
        e->top_nodes[4 * j + 3] = n3;
    }
    sc.nodes = e->nodes.data();
    sc.top_nodes = e->top_nodes.data();
    sc.prims = e->packed.data();
    sc.prim_ext = e->ext.data();
    sc.leaf_prim = e->leaf_prim.data();
    sc.leaf_mat = e->leaf_mat.data();
    sc.n_leaves = n;
    sc.n_internal = ni;
    sc.over_first = m;
    sc.over_count = n > leaf_size ? n - m : 0;
    sc.n_top = (int)queue.size();
    // max_top < 0: walk the 4-wide nodes (k_widen in loop form), as rtw_build_accel does by default
    if (max_top < 0 && n > leaf_size) {
        e->wide.resize((size_t)8 * ni);
        for (int i = 0; i < ni; i++) make_wide_node(e->nodes.data(), i, e->wide.data() + (size_t)8 * i);
        sc.wide = e->wide.data();
        sc.n_top = 0;
    }
    if (n <= leaf_size) sc.root = leaf_range_ref(0, n);
    else sc.root = sc.n_top > 0 ? RTW_REF_TOP : 0u;
    return e;
}

extern "C" void emu_destroy(Emu* e) { delete e; }

extern "C" int emu_n_internal(const Emu* e) { return e->sc.n_internal; }

extern "C" void emu_bvh_download(const Emu* e, unsigned long long* morton, int* leaf_prim, int* left, int* right,
                                 float* bounds, float* scene_box) {
    const int n = e->sc.n_leaves, ni = e->sc.n_internal;
    memcpy(morton, e->keys.data(), sizeof(unsigned long long) * n);
    memcpy(leaf_prim, e->leaf_prim.data(), sizeof(int) * n);
    for (int i = 0; i < ni; i++) {
        left[i] = e->left[i];
        right[i] = e->right[i];
        const float4 n0 = e->nodes[4 * i], n1 = e->nodes[4 * i + 1], n2 = e->nodes[4 * i + 2];
        float* b = bounds + (size_t)12 * i;
        b[0] = n0.x; b[1] = n0.y; b[2] = n0.z; b[3] = n0.w; b[4] = n1.x; b[5] = n1.y;
        b[6] = n1.z; b[7] = n1.w; b[8] = n2.x; b[9] = n2.y; b[10] = n2.z; b[11] = n2.w;
    }
    for (int i = 0; i < 3; i++) { scene_box[i] = e->scene_min[i]; scene_box[3 + i] = e->scene_max[i]; }
}

static RayW to_ray(const rtw_ray& a) {
    RayW r;
    r.o = f3(a.origin[0], a.origin[1], a.origin[2]);
    r.d = f3(a.dir[0], a.dir[1], a.dir[2]);
    r.tmin = a.tmin; r.tmax = a.tmax; r.time = a.time; r.gather = a.gather_time; r.seed = a.seed;
    return r;
}

extern "C" void emu_trace(const Emu* e, const rtw_ray* rays, int n, rtw_hit* hits, int* occluded,
                          unsigned long long* nodes_visited, unsigned long long* prims_tested) {
    TravStats st{0, 0};
    unsigned long long nv = 0, pt = 0;
    for (int i = 0; i < n; i++) {
        const RayW r = to_ray(rays[i]);
        st.nodes = st.prims = 0;
        if (hits) {
            float t; uint32_t leaf;
            const bool f = traverse_closest<MemGeneric, true>(e->sc, e->sc.top_nodes, r, t, leaf, &st);
            hits[i].t = f ? t : 0.f;
            hits[i].prim = f ? e->sc.leaf_prim[leaf] : -1;
        }
        if (occluded) occluded[i] = traverse_any<MemGeneric, true>(e->sc, e->sc.top_nodes, r, &st) ? 1 : 0;
        nv += st.nodes; pt += st.prims;
    }
    if (nodes_visited) *nodes_visited = nv;
    if (prims_tested) *prims_tested = pt;
}

// The wavefront loop of rtw_render_accum with every kernel replaced by a loop
// over its per-thread body.  accum: Nx*Ny float4 (not cleared).
extern "C" int emu_render(const Emu* e, int Nx, int Ny, int sample_first, int sample_stride, int sample_count,
                          int max_depth, int P, float* accum, unsigned long long* stats /* rays, shadow rays, iterations */) {
    const DScene& sc = e->sc;
    std::vector<SlotRec> rec(P);
    std::vector<float4> rad(P), sh_o(P), sh_d(P), sh_c(P);
    std::vector<uint32_t> queue((size_t)RTW_Q_SLOTS * P), counters(RTW_C_WORDS, 0);
    DWave w{};
    w.rec = rec.data(); w.rad = rad.data();
    w.queue = queue.data(); w.sh_o = sh_o.data(); w.sh_d = sh_d.data(); w.sh_c = sh_c.data();
    w.counters = counters.data(); w.accum = reinterpret_cast<float4*>(accum);
    w.P = P; w.Nx = Nx; w.Ny = Ny; w.max_depth = max_depth;
    w.sample_first = sample_first; w.sample_stride = sample_stride;
    w.total = (unsigned long long)Nx * Ny * sample_count;
    w.pixel_first = 0; w.pixel_count = (uint32_t)(Nx * Ny);
    unsigned long long next = 0, rays = 0, shrays = 0, iters = 0;
    if (max_depth == 0) {   // k_add_count
        for (int i = 0; i < Nx * Ny; i++) accum[4 * i + 3] += (float)sample_count;
        if (stats) { stats[0] = 0; stats[1] = 0; stats[2] = 0; }
        return 0;
    }
    uint32_t* regen = queue.data() + (size_t)RTW_Q_REGEN * P;
    for (int slot = 0; slot < P; slot++) {   // k_wave_init
        retire_slot(w, slot);
        regen[slot] = (uint32_t)slot;
    }
    counters[RTW_C_QUEUE0 + RTW_Q_REGEN] = (uint32_t)P;
    for (;;) {
        for (uint32_t i = 0; i < counters[RTW_C_QUEUE0 + RTW_Q_REGEN]; i++) {  // k_regen
            const uint32_t slot = regen[i];
            logic_flush(w, w.rad[slot]);
            const unsigned long long ord = next++;
            if (ord < w.total) generate_path(sc, w, slot, ord);
            else retire_slot(w, slot);
        }
        for (int c = 0; c < 8; c++) counters[c] = 0;
        iters++;
        for (int slot = 0; slot < P; slot++) {  // k_traverse
            const int cls = traverse_thread<MemGeneric>(sc, sc.top_nodes, w, slot);
            if (cls >= 0) queue[(size_t)cls * P + counters[cls]++] = slot;
        }
        for (int cls = 0; cls < RTW_Q_COUNT; cls++) {  // k_shade<cls>
            for (uint32_t i = 0; i < counters[cls]; i++) {
                const uint32_t slot = queue[(size_t)cls * P + i];
                ShadowOut sh;
                bool fin;
                switch (cls) {
                case RTW_Q_LAMBERTIAN: fin = shade_thread<RTW_Q_LAMBERTIAN, MemGeneric>(sc, w, slot, sh); break;
                case RTW_Q_METAL: fin = shade_thread<RTW_Q_METAL, MemGeneric>(sc, w, slot, sh); break;
                case RTW_Q_DIELECTRIC: fin = shade_thread<RTW_Q_DIELECTRIC, MemGeneric>(sc, w, slot, sh); break;
                case RTW_Q_ISOTROPIC: fin = shade_thread<RTW_Q_ISOTROPIC, MemGeneric>(sc, w, slot, sh); break;
                default: fin = shade_thread<RTW_Q_TERMINAL, MemGeneric>(sc, w, slot, sh); break;
                }
                if (fin) regen[counters[RTW_C_QUEUE0 + RTW_Q_REGEN]++] = slot;
                if (sh.emit) {
                    const uint32_t pos = counters[RTW_C_SHADOW]++;
                    sh_o[pos] = make_float4(sh.o.x, sh.o.y, sh.o.z, sh.tmax);
                    sh_d[pos] = make_float4(sh.d.x, sh.d.y, sh.d.z, u2f(slot));
                    sh_c[pos] = make_float4(sh.contrib.x, sh.contrib.y, sh.contrib.z, u2f(sh.seed));
                }
            }
            rays += counters[cls];
        }
        shrays += counters[RTW_C_SHADOW];
        for (uint32_t i = 0; i < counters[RTW_C_SHADOW]; i++) shadow_thread<MemGeneric>(sc, sc.top_nodes, w, i);  // k_shadow
        uint32_t traced = 0;
        for (int cls = 0; cls < RTW_Q_COUNT; cls++) traced += counters[cls];
        if (traced == 0) break;   // nothing in flight (the regen pass above flushed the last finished paths)
    }
    if (stats) { stats[0] = rays; stats[1] = shrays; stats[2] = iters; }
    return 0;
}

extern "C" void emu_camera_rays(const Emu* e, int Nx, int Ny, int sample, rtw_ray* out) {
    const int npix = Nx * Ny;
    DWave w{};
    w.P = npix; w.Nx = Nx; w.Ny = Ny; w.sample_first = sample; w.sample_stride = 0; w.total = npix;
    w.pixel_first = 0; w.pixel_count = (uint32_t)npix;
    for (int i = 0; i < npix; i++) {
        const NewPath np = make_path(e->sc, w, (unsigned long long)i);
        rtw_ray& r = out[i];
        r.origin[0] = np.ray.o.x; r.origin[1] = np.ray.o.y; r.origin[2] = np.ray.o.z;
        r.dir[0] = np.ray.d.x; r.dir[1] = np.ray.d.y; r.dir[2] = np.ray.d.z;
        r.tmin = 1e-6f; r.tmax = 1.e27f; r.time = np.ray.o.w; r.gather_time = np.ray.d.w;
        r.seed = f2u(np.st.thr.w); r.pad = 0;
    }
}

extern "C" void emu_texture_eval(const Emu* e, int tex, float u, float v, const float* p, float* out) {
    float3 c = texture_eval(e->sc, tex, u, v, f3(p[0], p[1], p[2]));
    out[0] = c.x; out[1] = c.y; out[2] = c.z;
}

extern "C" void emu_rng_stream(uint32_t seed0, uint32_t seed1, int n, uint32_t* tea_out, uint32_t* lcg_out,
                               float* rnd_out, uint32_t* xs_out, float* randf_out) {
    const uint32_t s = tea<64>(seed0, seed1);
    *tea_out = s;
    uint32_t a = s, b = s, c = s, d = s;
    for (int i = 0; i < n; i++) {
        lcg_out[i] = lcg(a); rnd_out[i] = rnd(b); xs_out[i] = xorshift32(c); randf_out[i] = randf(d);
    }
}

extern "C" void emu_math_eval(int fn, const float* x, const float* y, int n, float* out) {
    for (int i = 0; i < n; i++) {
        switch (fn) {
            case RTW_FN_SIN: out[i] = det_sinf(x[i]); break;
            case RTW_FN_COS: out[i] = det_cosf(x[i]); break;
            case RTW_FN_LOG: out[i] = det_logf(x[i]); break;
            case RTW_FN_POW5: out[i] = det_pow5f(x[i]); break;
            case RTW_FN_ATAN2: out[i] = det_atan2f(x[i], y[i]); break;
            default: out[i] = det_asinf(x[i]); break;
        }
    }
}
