// rtw_ctx.h — the opaque context behind the C-ABI (include/rtw.h).
// Owns every device allocation (the reference's Director owned them and leaked
// most per-primitive buffers, SURVEY.md §8b "Ownership").
#pragma once
#include <cuda_runtime.h>
#include <string>
#include <vector>
#include "../kernels/internal.h"

struct rtw_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string error;

    // ---- scene (device copies) ----
    bool scene_loaded = false;
    bool accel_built = false;
    int n_prims = 0, n_xforms = 0, n_mats = 0;
    uint32_t class_mask = 0;
    std::vector<void*> scene_allocs;
    rtw::BuildBuffers build{};
    float4* wide_nodes = nullptr;     // 8 float4 per internal node (owned by scene_allocs)
    rtw::DScene scene{};
    rtw::LaunchCfg cfg{};
    rtw::SmallScene small{};          // filled by rtw_build_accel when the scene is one leaf range of <= RTW_SMALL_MAX
    int n_internal = 0;
    int leaf_size = 1;
    float scene_min[3] = {0, 0, 0}, scene_max[3] = {0, 0, 0};

    // ---- wavefront pools ----
    // A frame is rendered by `n_waves` independent pools ("waves"), each with its own stream, its own share of
    // the samples (sample k of a pixel goes to wave k mod n_waves) and its own counters; their kernels run
    // concurrently, so an instruction-bound traversal launch of one wave overlaps a DRAM-bound shading launch
    // of another (rtw_render.cu).  Wave 0 uses `stream`.
    static constexpr int kMaxWaves = 4;
    struct Wave {
        std::vector<void*> allocs;
        rtw::DWave w{};
        int capacity = 0;
        int key_bits = 0;                   // option sort_rays: the number of bins the arrays were allocated for
        bool for_stream = false;            // what the pool was allocated for (streaming wavefront or slot pool)
        cudaStream_t stream = nullptr;
        uint32_t* h_counters = nullptr;     // pinned ring of counter snapshots
        cudaEvent_t ring_events[8] = {};
        cudaEvent_t ev_done = nullptr;
        cudaEvent_t ev_iter[2] = {};        // lockstep: recorded at the start of even / odd iterations
    };
    Wave waves[kMaxWaves];
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
    std::vector<cudaEvent_t> prof_events;  // 5 per sampled iteration (profile option; wave 0)

    // ---- last frame ----
    float4* d_accum = nullptr;          // linear sums of the last rtw_render
    float4* d_out = nullptr;
    int frame_pixels = 0;
    int frame_spp = 0;

    // ---- options ----
    long long opt_pool_paths = 0;    // 0 = auto (rtw_render.cu)
    int opt_top_nodes = 1024;
    int opt_wide = 1;                // LBVH scenes: walk the 4-wide nodes (lbvh.h make_wide_node); 0 = the two-child nodes
    int opt_leaf_size = 0;
    int opt_small = -1;              // -1 = auto, 0 = never, 1 = whenever the scene qualifies
    int opt_small_static = 1;        // small scenes: rects and spheres in the static table (layout.h SmallGroup)
    int opt_shade_ctas = 3;          // k_shade_all<kCtas>: 2, 3 or 4 resident CTAs per SM
    int opt_walker = -1;             // -1 = auto (rtw_build_accel), 0 = flat kernels, 1 = persistent walkers
    int opt_step_nodes = 4, opt_step_prims = 8, opt_min_active = 12;   // walkers: measured best with binned rays (tools/sweep_walk{3,4,5}.sh; round 1, slot order: 16 / 8 / 22)
    int opt_stream = -1;             // -1 = auto (rtw_build_accel), 0 = slot pool, 1 = streaming wavefront
    int opt_grid_pct = 100;          // share of the resident-CTA grids one context launches (two contexts on one GPU: 50)
    int opt_waves = 0;               // 0 = auto (rtw_render.cu), else 1..4 concurrent pools per frame
    int opt_stagger = 0;             // odd waves start half an iteration after their even neighbour (measured slower)
    int opt_lockstep = 0;            // iteration i of every wave starts when all waves have finished iteration i - 1 (measured slower)
    int opt_tail_paths = 131072;     // slot pool: below this many paths alive (none left to start) one launch runs them to their end; 0 = off
    int opt_node_min = 0, opt_prim_min = 0;   // walkers: warp-vote thresholds of the inner loops (0 = plain step)
    int opt_fuse_shadow = 0;         // 1 = by-value small scenes trace their NEE rays inside k_shade_all (no shadow queue, no k_shadow_small launch).
                                     // Measured on the B200, C2 (profiles/r02/r02c_fuse_shadow.md): 2656 Mpaths/s fused against 3057 with the queue —
                                     // the separate shadow launch is what the OTHER wave's DRAM-bound shading launch overlaps with; default off
    // LBVH walkers take their rays bin by bin — camera rays by direction (cube map), bounce rays by (origin cell, direction
    // octant): 1 = closest-hit rays, 2 = NEE rays too, 0 = slot order, -1 = auto: 1 on every walker scene.  Measured
    // (profiles/r02/sort_rays_r02f.md): scene 4 468 -> 592 Mpaths/s, scene 2 901 -> 975 (walkers instead of the flat kernels),
    // stress scenes of 2^14 / 2^17 / 2^20 primitives 1285 -> 1515, 528 -> 594, 229 -> 260.
    int opt_sort_rays = -1;
    int opt_sort_camera = 1;         // camera rays get direction bins of their own (6 x 256 x 256 on a cube map)
    int opt_sort_obits = 5, opt_sort_dbits = 1;   // bits per axis of the origin cell / of the direction bin (obits + dbits <= 6: 2^18 bins)
    float key_lo[3] = {0, 0, 0}, key_hi[3] = {0, 0, 0};   // bounds of the primitive centroids (rtw_build_accel): the cells of the ray keys
    int opt_seed_base = 0;           // rtw_render: index of the frame's first sample (SURVEY §8b seed_base)
    int opt_profile = 0;
    int opt_lag = 3;

    uint32_t* d_overflow = nullptr;  // DScene::overflow: traversal-stack pushes that found the stack full

    rtw_stats stats{};
};

namespace rtw {

inline int fail(rtw_ctx* c, int code, const std::string& msg) {
    if (c) c->error = msg;
    return code;
}
inline int cuda_fail(rtw_ctx* c, cudaError_t e, const char* what) {
    return fail(c, RTW_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

#define RTW_CUDA(c, call)                                             \
    do {                                                              \
        cudaError_t e__ = (call);                                     \
        if (e__ != cudaSuccess) return rtw::cuda_fail((c), e__, #call); \
    } while (0)

// After a render / trace call: a traversal that ran out of stack dropped a subtree — fail loudly instead of
// returning a wrong answer (the word is cleared for the next call).
inline int check_overflow(rtw_ctx* c, uint64_t* count_out) {
    uint32_t v = 0;
    if (!c->d_overflow) return RTW_OK;
    cudaError_t e = cudaMemcpy(&v, c->d_overflow, sizeof(v), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return cuda_fail(c, e, "cudaMemcpy(overflow word)");
    if (count_out) *count_out = v;
    if (v == 0) return RTW_OK;
    cudaMemset(c->d_overflow, 0, sizeof(v));
    return fail(c, RTW_ERR_STATE, "traversal stack overflow: " + std::to_string(v) + " pushes dropped (hierarchy deeper than the stack)");
}

template <class T>
inline cudaError_t dalloc(std::vector<void*>& pool, T** out, size_t count) {
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, sizeof(T) * (count ? count : 1));
    if (e == cudaSuccess) {
        pool.push_back(p);
        *out = reinterpret_cast<T*>(p);
    }
    return e;
}
inline void dfree_all(std::vector<void*>& pool) {
    for (void* p : pool) cudaFree(p);
    pool.clear();
}

}  // namespace rtw
