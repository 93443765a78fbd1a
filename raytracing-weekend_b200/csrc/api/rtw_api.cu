// rtw_api.cu — C-ABI: context lifetime, scene upload, LBVH build.
// (render entry points: rtw_render.cu; test hooks: rtw_hooks.cu)
//
// Replaces Director::init / createScene / destroy of the reference
// (Director.cpp:33-104, 951-969; Director.h:120-126): no OptiX context, modules,
// program groups, pipeline or SBT — the "ABI" towards the device is the DScene
// struct passed by value to every kernel.
#include <cstdlib>
#include <cstdio>
#include <algorithm>
#include "rtw_ctx.h"
#include "../device/lbvh.h"

#include <cstring>

using namespace rtw;

static std::string g_create_error;

extern "C" int rtw_abi_version(void) { return RTW_ABI_VERSION; }

extern "C" int rtw_device_count(int* count) {
    if (!count) return RTW_ERR_INVALID;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        *count = 0;
        g_create_error = std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e);
        cudaGetLastError();
        return RTW_ERR_NO_DEVICE;
    }
    *count = n;
    return RTW_OK;
}

extern "C" const char* rtw_last_error(const rtw_ctx* ctx) { return ctx ? ctx->error.c_str() : g_create_error.c_str(); }

extern "C" int rtw_create(int device, rtw_ctx** out) {
    if (!out) return RTW_ERR_INVALID;
    *out = nullptr;
    int n = 0;
    int rc = rtw_device_count(&n);
    if (rc != RTW_OK || n <= 0) {
        if (rc == RTW_OK) g_create_error = "no CUDA device visible (this library has no CPU fallback)";
        return RTW_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= n) {
        g_create_error = "device index out of range";
        return RTW_ERR_INVALID;
    }
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) {
        g_create_error = std::string("cudaSetDevice: ") + cudaGetErrorString(e);
        return RTW_ERR_CUDA;
    }
    rtw_ctx* c = new rtw_ctx();
    c->device = device;
    e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreate(&c->ev_begin);
    if (e == cudaSuccess) e = cudaEventCreate(&c->ev_end);
    if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&c->d_overflow), sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMemset(c->d_overflow, 0, sizeof(uint32_t));
    if (e != cudaSuccess) {
        g_create_error = std::string("context setup: ") + cudaGetErrorString(e);
        delete c;
        return RTW_ERR_CUDA;
    }
    *out = c;
    return RTW_OK;
}

static void free_frame(rtw_ctx* c) {
    if (c->d_accum) cudaFree(c->d_accum);
    if (c->d_out) cudaFree(c->d_out);
    c->d_accum = nullptr;
    c->d_out = nullptr;
    c->frame_pixels = 0;
}

extern "C" int rtw_destroy(rtw_ctx* c) {
    if (!c) return RTW_ERR_INVALID;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    dfree_all(c->scene_allocs);
    for (auto& wv : c->waves) {
        if (wv.stream) cudaStreamSynchronize(wv.stream);
        dfree_all(wv.allocs);
        if (wv.h_counters) cudaFreeHost(wv.h_counters);
        for (auto& e : wv.ring_events) if (e) cudaEventDestroy(e);
        if (wv.ev_done) cudaEventDestroy(wv.ev_done);
        for (auto& e : wv.ev_iter) if (e) cudaEventDestroy(e);
        if (wv.stream && wv.stream != c->stream) cudaStreamDestroy(wv.stream);
    }
    free_frame(c);
    for (auto& e : c->prof_events) if (e) cudaEventDestroy(e);
    if (c->ev_begin) cudaEventDestroy(c->ev_begin);
    if (c->ev_end) cudaEventDestroy(c->ev_end);
    if (c->d_overflow) cudaFree(c->d_overflow);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return RTW_OK;
}

extern "C" int rtw_set_option(rtw_ctx* c, const char* name, int64_t value) {
    if (!c || !name) return RTW_ERR_INVALID;
    if (!strcmp(name, "pool_paths")) {
        if (value != 0 && (value < 1024 || value > (1ll << 26))) return fail(c, RTW_ERR_INVALID, "pool_paths out of range [1024, 2^26] (0 = auto)");
        c->opt_pool_paths = value;
    } else if (!strcmp(name, "top_nodes")) {
        if (value < 0 || value > 3584) return fail(c, RTW_ERR_INVALID, "top_nodes out of range [0, 3584]");
        c->opt_top_nodes = (int)value;
    } else if (!strcmp(name, "wide")) {
        c->opt_wide = value ? 1 : 0;
    } else if (!strcmp(name, "leaf_size")) {
        if (value < 0 || value > RTW_MAX_LEAF_SIZE) return fail(c, RTW_ERR_INVALID, "leaf_size out of range [0, 32]");
        c->opt_leaf_size = (int)value;
    } else if (!strcmp(name, "step_nodes") || !strcmp(name, "step_prims") || !strcmp(name, "min_active")) {
        if (value < 1 || value > 1024) return fail(c, RTW_ERR_INVALID, "walk option out of range [1, 1024]");
        if (name[0] == 'm') c->opt_min_active = (int)(value > 32 ? 32 : value);
        else if (name[5] == 'n') c->opt_step_nodes = (int)value;
        else c->opt_step_prims = (int)value;
        c->cfg.step_nodes = c->opt_step_nodes; c->cfg.step_prims = c->opt_step_prims; c->cfg.min_active = c->opt_min_active;
    } else if (!strcmp(name, "small")) {
        if (value < -1 || value > 1) return fail(c, RTW_ERR_INVALID, "small out of range [-1, 1]");
        c->opt_small = (int)value;
    } else if (!strcmp(name, "small_static")) {
        c->opt_small_static = value ? 1 : 0;   // 0: every record of a small scene goes through the runs (test / comparison)
    } else if (!strcmp(name, "walker")) {
        if (value < -1 || value > 1) return fail(c, RTW_ERR_INVALID, "walker out of range [-1, 1]");
        c->opt_walker = (int)value;
    } else if (!strcmp(name, "shade_ctas")) {
        if (value < 2 || value > 4) return fail(c, RTW_ERR_INVALID, "shade_ctas out of range [2, 4]");
        c->opt_shade_ctas = (int)value;
    } else if (!strcmp(name, "stream")) {
        if (value < -1 || value > 1) return fail(c, RTW_ERR_INVALID, "stream out of range [-1, 1]");
        c->opt_stream = (int)value;
    } else if (!strcmp(name, "grid_pct")) {
        if (value < 1 || value > 100) return fail(c, RTW_ERR_INVALID, "grid_pct out of range [1, 100]");
        c->opt_grid_pct = (int)value;
    } else if (!strcmp(name, "waves")) {
        if (value < 0 || value > rtw_ctx::kMaxWaves) return fail(c, RTW_ERR_INVALID, "waves out of range [0, 4] (0 = auto)");
        c->opt_waves = (int)value;
    } else if (!strcmp(name, "node_min") || !strcmp(name, "prim_min")) {
        if (value < 0 || value > 32) return fail(c, RTW_ERR_INVALID, "vote threshold out of range [0, 32]");
        if (name[0] == 'n') c->opt_node_min = (int)value;
        else c->opt_prim_min = (int)value;
        c->cfg.node_min = c->opt_node_min; c->cfg.prim_min = c->opt_prim_min;
    } else if (!strcmp(name, "tail_paths")) {
        if (value < 0 || value > (1 << 22)) return fail(c, RTW_ERR_INVALID, "tail_paths out of range [0, 2^22]");
        c->opt_tail_paths = (int)value;
    } else if (!strcmp(name, "sort_rays")) {
        if (value < -1 || value > 2) return fail(c, RTW_ERR_INVALID, "sort_rays out of range [-1, 2]");
        c->opt_sort_rays = (int)value;
    } else if (!strcmp(name, "seed_base")) {
        if (value < 0 || value > 0x3fffffff) return fail(c, RTW_ERR_INVALID, "seed_base out of range [0, 2^30)");
        c->opt_seed_base = (int)value;
    } else if (!strcmp(name, "sort_camera")) {
        c->opt_sort_camera = value != 0;
    } else if (!strcmp(name, "sort_obits") || !strcmp(name, "sort_dbits")) {
        if (value < 1 || value > 5) return fail(c, RTW_ERR_INVALID, "sort_obits / sort_dbits out of range [1, 5]");
        if (name[5] == 'o') c->opt_sort_obits = (int)value;
        else c->opt_sort_dbits = (int)value;   // (obits + dbits <= 6 is enforced where the bins are sized: the direction bits give way)
    } else if (!strcmp(name, "fuse_shadow")) {
        if (value < -1 || value > 1) return fail(c, RTW_ERR_INVALID, "fuse_shadow out of range [-1, 1]");
        c->opt_fuse_shadow = (int)value;
        c->cfg.fuse_shadow = c->cfg.small && (value > 0);
        if (c->accel_built) { cudaError_t e = wavefront_configure(c->scene, &c->cfg); if (e != cudaSuccess) return cuda_fail(c, e, "wavefront_configure"); }
    } else if (!strcmp(name, "lockstep")) {
        c->opt_lockstep = value ? 1 : 0;
    } else if (!strcmp(name, "stagger")) {
        c->opt_stagger = value ? 1 : 0;
    } else if (!strcmp(name, "profile")) {
        c->opt_profile = value ? 1 : 0;
    } else if (!strcmp(name, "lag")) {
        if (value < 1 || value > 7) return fail(c, RTW_ERR_INVALID, "lag out of range [1, 7]");
        c->opt_lag = (int)value;
    } else {
        return fail(c, RTW_ERR_INVALID, std::string("unknown option ") + name);
    }
    return RTW_OK;
}

extern "C" int rtw_get_stats(rtw_ctx* c, rtw_stats* out) {
    if (!c || !out) return RTW_ERR_INVALID;
    *out = c->stats;
    return RTW_OK;
}

// ---------------------------------------------------------------------------
// Scene upload: flat tables of Director::initLaunchParams / createSBT
// (Director.cpp:483-553, 659-848) without the per-record cudaMalloc churn.
// ---------------------------------------------------------------------------
const char* rtw_internal_validate_desc(const rtw_scene_desc* d);   // host/scene_builder.cpp: every index and enumeration a kernel follows
static int validate_scene(rtw_ctx* c, const rtw_scene_desc* d) {
    if (const char* why = rtw_internal_validate_desc(d)) return fail(c, RTW_ERR_INVALID, why);
    return RTW_OK;
}

extern "C" int rtw_upload_scene(rtw_ctx* c, const rtw_scene_desc* d) {
    if (!c || !d) return RTW_ERR_INVALID;
    int rc = validate_scene(c, d);
    if (rc != RTW_OK) return rc;
    RTW_CUDA(c, cudaSetDevice(c->device));
    RTW_CUDA(c, cudaStreamSynchronize(c->stream));
    dfree_all(c->scene_allocs);
    c->scene_loaded = false;
    c->accel_built = false;
    c->scene = DScene{};
    c->scene.overflow = c->d_overflow;
    c->build = BuildBuffers{};
    auto& pool = c->scene_allocs;
    const int n = d->n_prims;
    c->n_prims = n;
    c->n_xforms = d->n_transforms;
    c->n_mats = d->n_materials;

    // raw arrays for the builder
    rtw_prim* d_prims = nullptr;
    rtw_transform* d_xf = nullptr;
    rtw_material* d_mats_raw = nullptr;
    RTW_CUDA(c, dalloc(pool, &d_prims, n));
    RTW_CUDA(c, cudaMemcpyAsync(d_prims, d->prims, sizeof(rtw_prim) * n, cudaMemcpyHostToDevice, c->stream));
    if (d->n_transforms > 0) {
        RTW_CUDA(c, dalloc(pool, &d_xf, d->n_transforms));
        RTW_CUDA(c, cudaMemcpyAsync(d_xf, d->transforms, sizeof(rtw_transform) * d->n_transforms, cudaMemcpyHostToDevice, c->stream));
    }
    RTW_CUDA(c, dalloc(pool, &d_mats_raw, d->n_materials));
    RTW_CUDA(c, cudaMemcpyAsync(d_mats_raw, d->materials, sizeof(rtw_material) * d->n_materials, cudaMemcpyHostToDevice, c->stream));
    c->build.prims = d_prims;
    c->build.xforms = d_xf;
    c->build.mats_raw = d_mats_raw;
    c->build.n = n;

    // shading tables in their float4 device layout
    DScene& sc = c->scene;
    std::vector<float4> xf4((size_t)6 * (d->n_transforms > 0 ? d->n_transforms : 1));
    for (int i = 0; i < d->n_transforms; i++) {
        const rtw_transform& t = d->transforms[i];
        for (int r = 0; r < 3; r++) {
            xf4[6 * i + r] = make_float4(t.w2o[4 * r], t.w2o[4 * r + 1], t.w2o[4 * r + 2], t.w2o[4 * r + 3]);
            xf4[6 * i + 3 + r] = make_float4(t.o2w[4 * r], t.o2w[4 * r + 1], t.o2w[4 * r + 2], t.o2w[4 * r + 3]);
        }
    }
    float4* d_xf4 = nullptr;
    RTW_CUDA(c, dalloc(pool, &d_xf4, xf4.size()));
    RTW_CUDA(c, cudaMemcpyAsync(d_xf4, xf4.data(), sizeof(float4) * xf4.size(), cudaMemcpyHostToDevice, c->stream));
    sc.xforms = d_xf4;

    std::vector<float4> mats(d->n_materials);
    for (int i = 0; i < d->n_materials; i++)
        mats[i] = make_float4(u2f((uint32_t)d->materials[i].type), u2f((uint32_t)d->materials[i].texture), d->materials[i].param, 0.f);
    float4* d_mats = nullptr;
    RTW_CUDA(c, dalloc(pool, &d_mats, mats.size()));
    RTW_CUDA(c, cudaMemcpyAsync(d_mats, mats.data(), sizeof(float4) * mats.size(), cudaMemcpyHostToDevice, c->stream));
    sc.mats = d_mats;

    std::vector<float4> texs((size_t)2 * (d->n_textures > 0 ? d->n_textures : 1));
    for (int i = 0; i < d->n_textures; i++) {
        const rtw_texture& t = d->textures[i];
        texs[2 * i] = make_float4(u2f((uint32_t)t.type), t.color[0], t.color[1], t.color[2]);
        const int idx = t.type == RTW_TEX_NOISE ? t.perlin : t.image;
        texs[2 * i + 1] = make_float4(u2f((uint32_t)t.odd), u2f((uint32_t)t.even), t.scale, u2f((uint32_t)idx));
    }
    float4* d_texs = nullptr;
    RTW_CUDA(c, dalloc(pool, &d_texs, texs.size()));
    RTW_CUDA(c, cudaMemcpyAsync(d_texs, texs.data(), sizeof(float4) * texs.size(), cudaMemcpyHostToDevice, c->stream));
    sc.texs = d_texs;

    if (d->n_perlin > 0) {
        std::vector<float4> vec((size_t)256 * d->n_perlin);
        std::vector<int> perm((size_t)768 * d->n_perlin);
        for (int t = 0; t < d->n_perlin; t++) {
            const rtw_perlin& p = d->perlin[t];
            for (int i = 0; i < 256; i++) {
                vec[256 * t + i] = make_float4(p.ranvec[i][0], p.ranvec[i][1], p.ranvec[i][2], 0.f);
                perm[768 * t + i] = p.perm_x[i];
                perm[768 * t + 256 + i] = p.perm_y[i];
                perm[768 * t + 512 + i] = p.perm_z[i];
            }
        }
        float4* d_vec = nullptr;
        int* d_perm = nullptr;
        RTW_CUDA(c, dalloc(pool, &d_vec, vec.size()));
        RTW_CUDA(c, dalloc(pool, &d_perm, perm.size()));
        RTW_CUDA(c, cudaMemcpyAsync(d_vec, vec.data(), sizeof(float4) * vec.size(), cudaMemcpyHostToDevice, c->stream));
        RTW_CUDA(c, cudaMemcpyAsync(d_perm, perm.data(), sizeof(int) * perm.size(), cudaMemcpyHostToDevice, c->stream));
        RTW_CUDA(c, cudaStreamSynchronize(c->stream));  // host vectors go out of scope
        sc.perlin_vec = d_vec;
        sc.perlin_perm = d_perm;
    }
    if (d->n_images > 0) {
        std::vector<DImage> imgs(d->n_images);
        for (int i = 0; i < d->n_images; i++) {
            const rtw_image& im = d->images[i];
            if (im.width <= 0 || im.height <= 0 || !im.rgba) return fail(c, RTW_ERR_INVALID, "bad image");
            uchar4* d_tex = nullptr;
            RTW_CUDA(c, dalloc(pool, &d_tex, (size_t)im.width * im.height));
            RTW_CUDA(c, cudaMemcpyAsync(d_tex, im.rgba, (size_t)im.width * im.height * 4, cudaMemcpyHostToDevice, c->stream));
            imgs[i].texels = d_tex;
            imgs[i].width = im.width;
            imgs[i].height = im.height;
        }
        DImage* d_imgs = nullptr;
        RTW_CUDA(c, dalloc(pool, &d_imgs, imgs.size()));
        RTW_CUDA(c, cudaMemcpyAsync(d_imgs, imgs.data(), sizeof(DImage) * imgs.size(), cudaMemcpyHostToDevice, c->stream));
        RTW_CUDA(c, cudaStreamSynchronize(c->stream));
        sc.images = d_imgs;
        sc.has_images = 1;
    }

    // flags, lights, camera
    c->class_mask = 1u << RTW_Q_TERMINAL;
    for (int i = 0; i < n; i++) {
        const int pt = d->prims[i].type;
        if (pt == RTW_PRIM_VOLUME_BOX || pt == RTW_PRIM_VOLUME_SPHERE) sc.has_media = 1;
        if (pt == RTW_PRIM_MOVING_SPHERE) sc.has_motion = 1;
        switch (d->materials[d->prims[i].material].type) {
        case RTW_MAT_LAMBERTIAN: c->class_mask |= 1u << RTW_Q_LAMBERTIAN; break;
        case RTW_MAT_METAL: c->class_mask |= 1u << RTW_Q_METAL; break;
        case RTW_MAT_DIELECTRIC: c->class_mask |= 1u << RTW_Q_DIELECTRIC; break;
        case RTW_MAT_ISOTROPIC: c->class_mask |= 1u << RTW_Q_ISOTROPIC; break;
        default: break;
        }
    }
    sc.n_lights = d->n_lights;
    sc.sky_light = d->sky_light;
    sc.has_light_rect = d->has_light_rect;
    sc.rect_axis = d->light_rect.axis;
    sc.rect_a0 = d->light_rect.a0; sc.rect_a1 = d->light_rect.a1;
    sc.rect_b0 = d->light_rect.b0; sc.rect_b1 = d->light_rect.b1;
    sc.rect_k = d->light_rect.k;
    for (int i = 0; i < d->n_lights; i++) {
        const rtw_light& l = d->lights[i];
        sc.lights[i].normal = f3(l.normal[0], l.normal[1], l.normal[2]);
        sc.lights[i].area = l.area;
        sc.lights[i].emission = f3(l.emission[0], l.emission[1], l.emission[2]);
        sc.lights[i].pad = 0.f;
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
    RTW_CUDA(c, cudaStreamSynchronize(c->stream));  // borrowed host arrays are free again
    c->scene_loaded = true;
    return RTW_OK;
}

// ---------------------------------------------------------------------------
// LBVH build (replaces all optixAccelBuild call sites)
// ---------------------------------------------------------------------------
extern "C" int rtw_build_accel(rtw_ctx* c) {
    if (!c) return RTW_ERR_INVALID;
    if (!c->scene_loaded) return fail(c, RTW_ERR_STATE, "rtw_build_accel before rtw_upload_scene");
    RTW_CUDA(c, cudaSetDevice(c->device));
    auto& pool = c->scene_allocs;
    BuildBuffers& b = c->build;
    const int n = b.n;
    const int ni = n > 1 ? n - 1 : 1;
    if (!b.box_mn) {  // first build for this scene: allocate
        RTW_CUDA(c, dalloc(pool, &b.box_mn, n));
        RTW_CUDA(c, dalloc(pool, &b.box_mx, n));
        RTW_CUDA(c, dalloc(pool, &b.bounds, 16));
        RTW_CUDA(c, dalloc(pool, &b.keys[0], n));
        RTW_CUDA(c, dalloc(pool, &b.keys[1], n));
        RTW_CUDA(c, dalloc(pool, &b.vals[0], n));
        RTW_CUDA(c, dalloc(pool, &b.vals[1], n));
        RTW_CUDA(c, dalloc(pool, &b.hist, sort_scratch_words(n)));
        RTW_CUDA(c, dalloc(pool, &b.parent, ni));
        RTW_CUDA(c, dalloc(pool, &b.leaf_parent, n));
        RTW_CUDA(c, dalloc(pool, &b.flags, ni));
        RTW_CUDA(c, dalloc(pool, &b.top_index, ni));
        RTW_CUDA(c, dalloc(pool, &b.leaf_mn, n));
        RTW_CUDA(c, dalloc(pool, &b.leaf_mx, n));
        RTW_CUDA(c, dalloc(pool, &b.prims_packed, (size_t)2 * n));
        RTW_CUDA(c, dalloc(pool, &b.prim_ext, n));
        RTW_CUDA(c, dalloc(pool, &b.leaf_prim, n));
        RTW_CUDA(c, dalloc(pool, &b.leaf_mat, n));
        RTW_CUDA(c, dalloc(pool, &b.nodes, (size_t)4 * ni));
        RTW_CUDA(c, dalloc(pool, &b.top_nodes, (size_t)4 * 3584));
        RTW_CUDA(c, dalloc(pool, &c->wide_nodes, (size_t)8 * ni));
    }
    b.max_top = c->opt_top_nodes;
    // leaf ranges: a scene of <= 16 primitives is ONE range (every lane of a warp walks the
    // same primitive list: no traversal divergence at all); larger scenes collapse subtrees
    // of <= 4 primitives, which trades a few extra primitive tests for fewer node visits.
    // (the walker rule explained at the traversal-variant block below; decided here because the leaf size follows it)
    const bool nee_stage_pre = c->scene.n_lights > 0 && c->scene.has_light_rect && (c->class_mask & (1u << RTW_Q_LAMBERTIAN));
    const bool walker_pre = c->opt_walker >= 0 ? c->opt_walker != 0 : (n > 262144 || (nee_stage_pre && n > RTW_SMALL_MAX));
    // Plain LBVH leaves (leaf size 1) for every scene that is walked through the hierarchy: under a 4-wide node a leaf
    // range saves no fetch, and the primitive tests of a range are what diverges (mixed kinds; lanes with 1 and lanes
    // with 4 primitives).  Leaf size 1 against round 1's 4 (tuned on the two-child walk), profiles/r02/lbvh_r02.md,
    // tools/sweep_walk{3,4,5}.sh, tools/gpu_auto_check{2,3}.sh: walkers +8 % on scene 4, +30 % / +45 % on the stress
    // scenes of 2^20 / 2^17 primitives; streaming wavefront +11 % on the Book-1 scene, +34 % / +66 % at 2048 / 16 384.
    b.leaf_size = c->opt_leaf_size > 0 ? c->opt_leaf_size : (n <= 16 ? 16 : 1);
    c->leaf_size = b.leaf_size;
    const bool wide = c->opt_wide != 0 && n > b.leaf_size;
    b.wide = wide ? c->wide_nodes : nullptr;
    uint64_t launches = 0;
    RTW_CUDA(c, cudaEventRecord(c->ev_begin, c->stream));
    RTW_CUDA(c, lbvh_build(c->stream, b, &launches));
    RTW_CUDA(c, cudaEventRecord(c->ev_end, c->stream));
    uint32_t hb[16];   // lbvh.cu k_bounds_init: scene box, n_top, centroid bounds, primitives flagged oversize
    RTW_CUDA(c, cudaMemcpyAsync(hb, b.bounds, sizeof(hb), cudaMemcpyDeviceToHost, c->stream));
    RTW_CUDA(c, cudaStreamSynchronize(c->stream));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, c->ev_begin, c->ev_end);
    c->stats.build_ms = ms;
    auto dec = [](uint32_t u) { return u2f((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u); };
    for (int i = 0; i < 3; i++) {
        c->scene_min[i] = dec(hb[i]);
        c->scene_max[i] = dec(hb[3 + i]);
        c->key_lo[i] = dec(hb[8 + i]);
        c->key_hi[i] = dec(hb[11 + i]);
    }
    DScene& sc = c->scene;
    sc.nodes = b.nodes;
    sc.top_nodes = b.top_nodes;
    sc.prims = b.prims_packed;
    sc.prim_ext = b.prim_ext;
    sc.leaf_prim = b.leaf_prim;
    sc.leaf_mat = b.leaf_mat;
    // oversize primitives (device/lbvh.h): the hierarchy covers the first n_tree leaves, every ray tests the rest directly
    const int n_tree = tree_leaf_count(n, (int)hb[14]);
    const int ni_tree = n_tree > 1 ? n_tree - 1 : 1;
    sc.n_leaves = n;
    sc.n_internal = ni_tree;
    sc.over_first = n_tree;
    sc.over_count = n > b.leaf_size ? n - n_tree : 0;   // a one-range scene walks all its leaves anyway
    sc.n_top = (int)hb[6];
    sc.wide = b.wide;
    if (wide) sc.n_top = 0;   // the wide walk reads its 128-byte records through L1 (no shared-memory staging)
    if (n <= b.leaf_size) sc.root = leaf_range_ref(0, n);
    else sc.root = sc.n_top > 0 ? (RTW_REF_TOP | 0u) : 0u;
    c->n_internal = ni_tree;
    // Traversal variant.  Persistent walkers (dynamic ray fetching) win on deep hierarchies (round 1, tools/sweep_walk.sh:
    // Book-2 final scene, 3410 primitives, k_traverse 0.77 -> 0.60 ms per 2 M rays); on the 520-528 primitive scenes
    // they lost ~10 % to the flat kernels until the rays were taken bin by bin (round 2, option sort_rays,
    // profiles/r02/sort_rays_r02f.md): with the binning the walkers also win on the Book-2 spheres scene (520 primitives,
    // NEE: 975 against 901 Mpaths/s), while the Book-1 scene (528 primitives, no NEE) stays with the streaming
    // wavefront (1821 against 1493 with plain leaves); on the sky-lit stress scenes streaming leads up to 2^17 primitives
    // (4297 against 3604 at 2048, 2533 against 2195 at 16 384, 984 against 842 at 2^17) and falls behind at 2^20 (281
    // against 334).  Auto: walkers for every LBVH scene with an NEE stage, and above 2^18 primitives.
    c->cfg.walker = walker_pre;   // (decided before the build: the leaf size follows it)
    // Small scenes (one leaf range, <= RTW_SMALL_MAX leaves and transforms): the leaf records go to
    // the traversal kernels by value.
    c->cfg.small = nullptr;
    c->cfg.shade_ctas = c->opt_shade_ctas;
    if (c->opt_small != 0 && !(c->opt_walker == 1) && (sc.root & RTW_REF_LEAF) && n <= RTW_SMALL_MAX && c->n_xforms <= RTW_SMALL_MAX) {
        SmallScene& ss = c->small;
        memset(&ss, 0, sizeof(ss));
        ss.n = n;
        float4 xf[6 * RTW_SMALL_MAX], pk[2 * RTW_SMALL_MAX];
        float2 ext[RTW_SMALL_MAX];
        int32_t ids[RTW_SMALL_MAX];
        RTW_CUDA(c, cudaMemcpyAsync(pk, b.prims_packed, sizeof(float4) * 2 * n, cudaMemcpyDeviceToHost, c->stream));
        RTW_CUDA(c, cudaMemcpyAsync(ext, b.prim_ext, sizeof(float2) * n, cudaMemcpyDeviceToHost, c->stream));
        RTW_CUDA(c, cudaMemcpyAsync(ids, b.leaf_prim, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, c->stream));
        if (c->n_xforms > 0)
            RTW_CUDA(c, cudaMemcpyAsync(xf, sc.xforms, sizeof(float4) * 6 * c->n_xforms, cudaMemcpyDeviceToHost, c->stream));
        RTW_CUDA(c, cudaStreamSynchronize(c->stream));
        // Rects and static spheres go to the static table (layout.h SmallGroup) while its shape has
        // room: group 0 = untransformed, groups 1.. = the transforms in order of first use.  The rest
        // stays in records ordered by (transform, kind), cut into runs (layout.h SmallRun).
        int group_xf[RTW_SS_GROUPS];
        ss.n_groups = 1;
        group_xf[0] = -1;
        int order[RTW_SMALL_MAX], key[RTW_SMALL_MAX], n_left = 0;
        for (int i = 0; i < n; i++) {
            const uint32_t meta = f2u(pk[2 * i + 1].w);
            const int kind = meta_type(meta);
            const int xfi = kind == RTW_PRIM_MOVING_SPHERE ? -1 : meta_xform(meta);
            const bool rect = kind == RTW_PRIM_RECT_X || kind == RTW_PRIM_RECT_Y || kind == RTW_PRIM_RECT_Z;
            bool placed = false;
            if (c->opt_small_static && (rect || kind == RTW_PRIM_SPHERE)) {
                int g = 0;
                while (g < ss.n_groups && group_xf[g] != xfi) g++;
                if (g == ss.n_groups && g < RTW_SS_GROUPS) {
                    group_xf[g] = xfi;
                    for (int k = 0; k < 3; k++) ss.group[g].w2o[k] = xf[6 * xfi + k];
                    ss.n_groups++;
                }
                if (g < ss.n_groups) {
                    SmallGroup& sg = ss.group[g];
                    if (rect) {
                        const int axis = kind - RTW_PRIM_RECT_X;
                        if (sg.n_rect[axis] < RTW_SS_RECTS) {
                            SmallRect& sr = sg.rect[axis][sg.n_rect[axis]++];
                            sr.a = pk[2 * i];
                            sr.k = pk[2 * i + 1].x;
                            sr.prim_id = ids[i];
                            sr.leaf = i;
                            placed = true;
                        }
                    } else if (sg.n_sphere < RTW_SS_SPHERES) {
                        SmallSphere& sp = sg.sphere[sg.n_sphere++];
                        sp.c = pk[2 * i];
                        sp.prim_id = ids[i];
                        sp.leaf = i;
                        placed = true;
                    }
                }
            }
            if (!placed) {
                key[i] = (xfi + 1) * 8 + kind;
                order[n_left++] = i;
            }
        }
        std::stable_sort(order, order + n_left, [&](int x, int y) { return key[x] < key[y]; });
        ss.n_runs = 0;
        for (int k = 0; k < n_left; k++) {
            const int i = order[k];
            ss.leaf[k].a = pk[2 * i];
            ss.leaf[k].b = pk[2 * i + 1];
            ss.leaf[k].ext = ext[i];
            ss.leaf[k].prim_id = ids[i];
            ss.leaf[k].leaf = i;
            if (k == 0 || key[i] != key[order[k - 1]]) {
                SmallRun& run = ss.run[ss.n_runs++];
                run.first = k;
                run.count = 0;
                run.kind = key[i] & 7;
                run.xf = (key[i] >> 3) - 1;
            }
            ss.run[ss.n_runs - 1].count++;
        }
        for (int j = 0; j < c->n_xforms; j++)
            for (int k = 0; k < 3; k++) ss.w2o[3 * j + k] = xf[6 * j + k];
        c->cfg.small = &c->small;
    }
    c->cfg.fuse_shadow = c->cfg.small && c->opt_fuse_shadow > 0;
    RTW_CUDA(c, wavefront_configure(sc, &c->cfg));
    if (c->opt_grid_pct < 100) {
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
        auto share = [&](int& g) { g = std::max(sms, (g * c->opt_grid_pct / 100 + sms / 2) / sms * sms); };
        share(c->cfg.grid_logic); share(c->cfg.grid_traverse); share(c->cfg.grid_shade); share(c->cfg.grid_shadow);
    }
    c->cfg.class_mask = c->class_mask;
    c->cfg.step_nodes = c->opt_step_nodes; c->cfg.step_prims = c->opt_step_prims; c->cfg.min_active = c->opt_min_active;
    c->cfg.node_min = c->opt_node_min; c->cfg.prim_min = c->opt_prim_min;
    if (sc.root & RTW_REF_LEAF) {   // one leaf range: every ray does the same walk, nothing to rebalance
        c->cfg.step_nodes = 1; c->cfg.step_prims = RTW_MAX_LEAF_SIZE; c->cfg.min_active = 0;   // static mode
    }
    c->cfg.has_shadow = (sc.n_lights > 0 && sc.has_light_rect && (c->class_mask & (1u << RTW_Q_LAMBERTIAN)));
    // Streaming wavefront (kernels/stream.cu).  Measured on the B200 (tools/gpu_stream_scenes.sh): +26 % on the
    // Book-1 scene (528 primitives, sky light, no NEE: 1398 -> 1763 Mpaths/s), where one LBVH walk per vertex
    // rides in the shading kernel; slower where a vertex also traces a shadow ray through the LBVH (scenes 2, 4:
    // two divergent walks at 2-3 CTAs per SM) and on the by-value small scenes (-5 %: the fused kernel is 170 KB
    // of code against a 32 KB instruction cache).  Auto = LBVH scenes without an NEE stage, flat traversal.
    c->cfg.stream = c->opt_stream >= 0 ? c->opt_stream != 0 : (!c->cfg.small && !c->cfg.has_shadow && !c->cfg.walker);
    RTW_CUDA(c, stream_configure(sc, &c->cfg));
    c->accel_built = true;
    return RTW_OK;
}
