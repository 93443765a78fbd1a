// rtw_hooks.cu — tier-1 test hooks of the C-ABI: closest /
// any hit on a caller-supplied ray set, RNG streams, camera rays, LBVH download,
// the builder's radix sort.  Host buffers in, host buffers out.
#include "rtw_ctx.h"

#include <vector>

using namespace rtw;

namespace {
struct Tmp {  // scoped device allocations
    std::vector<void*> pool;
    ~Tmp() { dfree_all(pool); }
};
}  // namespace

static int trace_common(rtw_ctx* c, const rtw_ray* rays, int n, rtw_hit* hits, int32_t* occ) {
    if (!c || !rays || n < 0) return RTW_ERR_INVALID;
    if (!c->accel_built) return fail(c, RTW_ERR_STATE, "trace before rtw_build_accel");
    if (n == 0) return RTW_OK;
    RTW_CUDA(c, cudaSetDevice(c->device));
    Tmp t;
    rtw_ray* d_rays = nullptr;
    rtw_hit* d_hits = nullptr;
    int32_t* d_occ = nullptr;
    RTW_CUDA(c, dalloc(t.pool, &d_rays, n));
    if (hits) RTW_CUDA(c, dalloc(t.pool, &d_hits, n));
    if (occ) RTW_CUDA(c, dalloc(t.pool, &d_occ, n));
    uint32_t* d_work = nullptr;
    RTW_CUDA(c, dalloc(t.pool, &d_work, 2));
    RTW_CUDA(c, cudaMemcpyAsync(d_rays, rays, sizeof(rtw_ray) * n, cudaMemcpyHostToDevice, c->stream));
    uint64_t l = 0;
    RTW_CUDA(c, launch_trace(c->stream, c->scene, c->cfg, d_rays, n, d_hits, d_occ, d_work, &l));
    if (hits) RTW_CUDA(c, cudaMemcpyAsync(hits, d_hits, sizeof(rtw_hit) * n, cudaMemcpyDeviceToHost, c->stream));
    if (occ) RTW_CUDA(c, cudaMemcpyAsync(occ, d_occ, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, c->stream));
    RTW_CUDA(c, cudaStreamSynchronize(c->stream));
    return check_overflow(c, nullptr);
}

extern "C" int rtw_trace_closest(rtw_ctx* c, const rtw_ray* rays, int n, rtw_hit* hits) {
    if (!hits) return RTW_ERR_INVALID;
    return trace_common(c, rays, n, hits, nullptr);
}
extern "C" int rtw_trace_any(rtw_ctx* c, const rtw_ray* rays, int n, int32_t* occluded) {
    if (!occluded) return RTW_ERR_INVALID;
    return trace_common(c, rays, n, nullptr, occluded);
}

extern "C" int rtw_rng_stream(rtw_ctx* c, uint32_t seed0, uint32_t seed1, int n, uint32_t* tea_out, uint32_t* lcg_out,
                              float* rnd_out, uint32_t* xs_out, float* randf_out) {
    if (!c || n < 0) return RTW_ERR_INVALID;
    RTW_CUDA(c, cudaSetDevice(c->device));
    Tmp t;
    uint32_t *d_tea = nullptr, *d_lcg = nullptr, *d_xs = nullptr;
    float *d_rnd = nullptr, *d_randf = nullptr;
    RTW_CUDA(c, dalloc(t.pool, &d_tea, 1));
    RTW_CUDA(c, dalloc(t.pool, &d_lcg, n));
    RTW_CUDA(c, dalloc(t.pool, &d_xs, n));
    RTW_CUDA(c, dalloc(t.pool, &d_rnd, n));
    RTW_CUDA(c, dalloc(t.pool, &d_randf, n));
    RTW_CUDA(c, launch_rng_stream(c->stream, seed0, seed1, n, d_tea, d_lcg, d_rnd, d_xs, d_randf));
    RTW_CUDA(c, cudaStreamSynchronize(c->stream));
    if (tea_out) RTW_CUDA(c, cudaMemcpy(tea_out, d_tea, 4, cudaMemcpyDeviceToHost));
    if (n > 0) {
        if (lcg_out) RTW_CUDA(c, cudaMemcpy(lcg_out, d_lcg, 4 * (size_t)n, cudaMemcpyDeviceToHost));
        if (rnd_out) RTW_CUDA(c, cudaMemcpy(rnd_out, d_rnd, 4 * (size_t)n, cudaMemcpyDeviceToHost));
        if (xs_out) RTW_CUDA(c, cudaMemcpy(xs_out, d_xs, 4 * (size_t)n, cudaMemcpyDeviceToHost));
        if (randf_out) RTW_CUDA(c, cudaMemcpy(randf_out, d_randf, 4 * (size_t)n, cudaMemcpyDeviceToHost));
    }
    return RTW_OK;
}

extern "C" int rtw_math_eval(rtw_ctx* c, int fn, const float* x, const float* y, int n, float* out) {
    if (!c || fn < 0 || fn >= RTW_FN_COUNT || !x || !out || n < 0 || (fn == RTW_FN_ATAN2 && !y)) return RTW_ERR_INVALID;
    if (n == 0) return RTW_OK;
    RTW_CUDA(c, cudaSetDevice(c->device));
    Tmp t;
    float *d_x = nullptr, *d_y = nullptr, *d_o = nullptr;
    RTW_CUDA(c, dalloc(t.pool, &d_x, n));
    RTW_CUDA(c, dalloc(t.pool, &d_o, n));
    RTW_CUDA(c, cudaMemcpyAsync(d_x, x, sizeof(float) * n, cudaMemcpyHostToDevice, c->stream));
    if (fn == RTW_FN_ATAN2) {
        RTW_CUDA(c, dalloc(t.pool, &d_y, n));
        RTW_CUDA(c, cudaMemcpyAsync(d_y, y, sizeof(float) * n, cudaMemcpyHostToDevice, c->stream));
    }
    RTW_CUDA(c, launch_math_eval(c->stream, fn, d_x, d_y, n, d_o));
    RTW_CUDA(c, cudaMemcpyAsync(out, d_o, sizeof(float) * n, cudaMemcpyDeviceToHost, c->stream));
    RTW_CUDA(c, cudaStreamSynchronize(c->stream));
    return RTW_OK;
}

extern "C" int rtw_camera_rays(rtw_ctx* c, int Nx, int Ny, int sample, rtw_ray* rays_out) {
    if (!c || Nx <= 0 || Ny <= 0 || sample < 0 || !rays_out) return RTW_ERR_INVALID;
    if (!c->scene_loaded) return fail(c, RTW_ERR_STATE, "camera rays before rtw_upload_scene");
    RTW_CUDA(c, cudaSetDevice(c->device));
    const int npix = Nx * Ny;
    Tmp t;
    DWave w{};
    rtw_ray* d_rays = nullptr;
    RTW_CUDA(c, dalloc(t.pool, &d_rays, npix));
    w.P = npix;
    w.Nx = Nx;
    w.Ny = Ny;
    w.total = (unsigned long long)npix;
    RTW_CUDA(c, launch_camera_rays(c->stream, c->scene, w, sample, d_rays));
    RTW_CUDA(c, cudaMemcpyAsync(rays_out, d_rays, sizeof(rtw_ray) * npix, cudaMemcpyDeviceToHost, c->stream));
    RTW_CUDA(c, cudaStreamSynchronize(c->stream));
    return RTW_OK;
}

extern "C" int rtw_bvh_info_get(rtw_ctx* c, rtw_bvh_info* info) {
    if (!c || !info) return RTW_ERR_INVALID;
    if (!c->accel_built) return fail(c, RTW_ERR_STATE, "bvh info before rtw_build_accel");
    info->n_leaves = c->scene.n_leaves;
    info->n_internal = c->scene.n_internal;
    info->root = 0;
    info->n_top = c->scene.n_top;
    for (int i = 0; i < 3; i++) {
        info->scene_min[i] = c->scene_min[i];
        info->scene_max[i] = c->scene_max[i];
    }
    return RTW_OK;
}

extern "C" int rtw_bvh_download(rtw_ctx* c, uint64_t* morton, int32_t* leaf_prim, int32_t* left, int32_t* right,
                                float* bounds) {
    if (!c) return RTW_ERR_INVALID;
    if (!c->accel_built) return fail(c, RTW_ERR_STATE, "bvh download before rtw_build_accel");
    RTW_CUDA(c, cudaSetDevice(c->device));
    RTW_CUDA(c, cudaStreamSynchronize(c->stream));
    const int n = c->scene.n_leaves, ni = c->scene.n_internal;
    if (morton) RTW_CUDA(c, cudaMemcpy(morton, c->build.keys[0], sizeof(uint64_t) * n, cudaMemcpyDeviceToHost));
    if (leaf_prim) RTW_CUDA(c, cudaMemcpy(leaf_prim, c->build.leaf_prim, sizeof(int32_t) * n, cudaMemcpyDeviceToHost));
    if (left || right || bounds) {
        std::vector<float4> nodes((size_t)4 * ni);
        RTW_CUDA(c, cudaMemcpy(nodes.data(), c->build.nodes, sizeof(float4) * nodes.size(), cudaMemcpyDeviceToHost));
        auto dl = [](uint32_t ref) -> int32_t { return (ref & RTW_REF_LEAF) ? ~(int32_t)(ref & RTW_REF_MASK) : (int32_t)ref; };
        for (int i = 0; i < ni; i++) {
            const float4 n0 = nodes[4 * i], n1 = nodes[4 * i + 1], n2 = nodes[4 * i + 2], n3 = nodes[4 * i + 3];
            if (left) left[i] = dl(f2u(n3.z));     // the Karras hierarchy (n3.xy are the collapsed traversal refs)
            if (right) right[i] = dl(f2u(n3.w));
            if (bounds) {
                float* b = bounds + (size_t)12 * i;
                b[0] = n0.x; b[1] = n0.y; b[2] = n0.z; b[3] = n0.w; b[4] = n1.x; b[5] = n1.y;
                b[6] = n1.z; b[7] = n1.w; b[8] = n2.x; b[9] = n2.y; b[10] = n2.z; b[11] = n2.w;
            }
        }
    }
    return RTW_OK;
}

extern "C" int rtw_sort_pairs(rtw_ctx* c, uint64_t* keys, uint32_t* values, int n) {
    if (!c || !keys || !values || n < 0) return RTW_ERR_INVALID;
    if (n == 0) return RTW_OK;
    RTW_CUDA(c, cudaSetDevice(c->device));
    Tmp t;
    unsigned long long* k[2] = {nullptr, nullptr};
    uint32_t* v[2] = {nullptr, nullptr};
    uint32_t* hist = nullptr;
    RTW_CUDA(c, dalloc(t.pool, &k[0], n));
    RTW_CUDA(c, dalloc(t.pool, &k[1], n));
    RTW_CUDA(c, dalloc(t.pool, &v[0], n));
    RTW_CUDA(c, dalloc(t.pool, &v[1], n));
    RTW_CUDA(c, dalloc(t.pool, &hist, sort_scratch_words(n)));
    RTW_CUDA(c, cudaMemcpyAsync(k[0], keys, sizeof(uint64_t) * n, cudaMemcpyHostToDevice, c->stream));
    RTW_CUDA(c, cudaMemcpyAsync(v[0], values, sizeof(uint32_t) * n, cudaMemcpyHostToDevice, c->stream));
    RTW_CUDA(c, sort_pairs(c->stream, k, v, n, hist, nullptr));
    RTW_CUDA(c, cudaMemcpyAsync(keys, k[0], sizeof(uint64_t) * n, cudaMemcpyDeviceToHost, c->stream));
    RTW_CUDA(c, cudaMemcpyAsync(values, v[0], sizeof(uint32_t) * n, cudaMemcpyDeviceToHost, c->stream));
    RTW_CUDA(c, cudaStreamSynchronize(c->stream));
    return RTW_OK;
}

// ---- GPU-side scene flattening / texture baking (kernels/bake.cu) ----
extern "C" int rtw_bake_transforms(rtw_ctx* c, const float* o2w_rows, int n, rtw_transform* out) {
    if (!c || !o2w_rows || !out || n < 0) return RTW_ERR_INVALID;
    if (n == 0) return RTW_OK;
    RTW_CUDA(c, cudaSetDevice(c->device));
    Tmp t;
    float* d_in = nullptr;
    rtw_transform* d_out = nullptr;
    RTW_CUDA(c, dalloc(t.pool, &d_in, (size_t)12 * n));
    RTW_CUDA(c, dalloc(t.pool, &d_out, n));
    RTW_CUDA(c, cudaMemcpyAsync(d_in, o2w_rows, sizeof(float) * 12 * n, cudaMemcpyHostToDevice, c->stream));
    RTW_CUDA(c, launch_invert_transforms(c->stream, d_in, n, d_out));
    RTW_CUDA(c, cudaMemcpyAsync(out, d_out, sizeof(rtw_transform) * n, cudaMemcpyDeviceToHost, c->stream));
    RTW_CUDA(c, cudaStreamSynchronize(c->stream));
    return RTW_OK;
}
extern "C" int rtw_bake_perlin(rtw_ctx* c, uint32_t mt_seed, int first_table, int n_tables, rtw_perlin* out) {
    if (!c || !out || first_table < 0 || n_tables < 0 || n_tables > 4096) return RTW_ERR_INVALID;
    if (n_tables == 0) return RTW_OK;
    RTW_CUDA(c, cudaSetDevice(c->device));
    Tmp t;
    float* d_draws = nullptr;
    rtw_perlin* d_out = nullptr;
    RTW_CUDA(c, dalloc(t.pool, &d_draws, (size_t)1533 * n_tables));
    RTW_CUDA(c, dalloc(t.pool, &d_out, n_tables));
    RTW_CUDA(c, launch_perlin_tables(c->stream, mt_seed, first_table, n_tables, d_draws, d_out));
    RTW_CUDA(c, cudaMemcpyAsync(out, d_out, sizeof(rtw_perlin) * n_tables, cudaMemcpyDeviceToHost, c->stream));
    RTW_CUDA(c, cudaStreamSynchronize(c->stream));
    return RTW_OK;
}
extern "C" int rtw_bake_image(rtw_ctx* c, const uint8_t* pixels, int w, int h, int channels, uint8_t* rgba_out) {
    if (!c || !pixels || !rgba_out || w <= 0 || h <= 0 || (channels != 3 && channels != 4) || (long long)w * h > 0x7fffffffll) return RTW_ERR_INVALID;
    RTW_CUDA(c, cudaSetDevice(c->device));
    Tmp t;
    uint8_t* d_in = nullptr;
    uchar4* d_out = nullptr;
    RTW_CUDA(c, dalloc(t.pool, &d_in, (size_t)w * h * channels));
    RTW_CUDA(c, dalloc(t.pool, &d_out, (size_t)w * h));
    RTW_CUDA(c, cudaMemcpyAsync(d_in, pixels, (size_t)w * h * channels, cudaMemcpyHostToDevice, c->stream));
    RTW_CUDA(c, launch_flip_rgba(c->stream, d_in, w, h, channels, d_out));
    RTW_CUDA(c, cudaMemcpyAsync(rgba_out, d_out, (size_t)w * h * 4, cudaMemcpyDeviceToHost, c->stream));
    RTW_CUDA(c, cudaStreamSynchronize(c->stream));
    return RTW_OK;
}
