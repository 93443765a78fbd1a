// internal.h — launch functions shared between the kernel translation units and
// the C-ABI implementation (api/rtw_api.cu).  Not part of the public boundary.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../device/layout.h"

namespace rtw {

// ---- radix sort (kernels/sort.cu) -----------------------------------------
// Stable LSD radix sort of (64-bit key, 32-bit value) pairs, 8 passes of 8 bits.
// keys/vals are ping-pong buffers of n elements each; the sorted result ends in
// keys[0]/vals[0].  scratch must hold sort_scratch_words(n) uint32 (digit totals per pass, per-tile digit counts);
// sort_pairs clears what it needs itself.
size_t sort_scratch_words(int n);
// passes: the low `passes` 8-bit digits are sorted (8 = whole keys); after an odd number the result is in keys[1]/vals[1].
cudaError_t sort_pairs(cudaStream_t s, unsigned long long* keys[2], uint32_t* vals[2], int n, uint32_t* scratch,
                       uint64_t* launches, int passes = 8);


// ---- LBVH build (kernels/lbvh.cu) -------------------------------------------
struct BuildBuffers {
    // inputs (device copies of the C-ABI scene arrays)
    const rtw_prim* prims;          // n
    const rtw_transform* xforms;    // may be null
    const rtw_material* mats_raw;   // for the material type in the leaf meta word
    int n;
    // scratch
    float4* box_mn;                 // n, prim order
    float4* box_mx;
    uint32_t* bounds;               // 6 order-preserving encoded floats + n_top word
    unsigned long long* keys[2];
    uint32_t* vals[2];
    uint32_t* hist;
    int* parent;                    // n_internal
    int* leaf_parent;               // n
    int* flags;                     // n_internal
    int* top_index;                 // n_internal
    // outputs
    float4* leaf_mn;                // n, leaf order
    float4* leaf_mx;
    float4* prims_packed;           // 2n
    float2* prim_ext;               // n
    int* leaf_prim;                 // n
    int* leaf_mat;                  // n
    float4* nodes;                  // 4 * n_internal
    float4* top_nodes;              // 4 * max_top
    float4* wide;                   // 8 * n_internal: 4-wide nodes (lbvh.h make_wide_node); built when not null
    int max_top;
    int leaf_size;                  // subtrees of <= leaf_size primitives become one leaf range (1 = plain LBVH)
};
cudaError_t lbvh_build(cudaStream_t s, const BuildBuffers& b, uint64_t* launches);

// ---- wavefront (kernels/wavefront.cu) --------------------------------------
struct LaunchCfg {
    int grid_logic, grid_traverse, grid_shade, grid_shadow;
    int block;
    size_t smem_top;        // bytes of shared memory for the staged top nodes
    uint32_t class_mask;    // which shading queues the scene can populate
    bool has_shadow;        // scene has lights → NEE stage runs
    bool fuse_shadow;       // small scenes: the shading kernel traces its NEE rays in place (no shadow queue, no k_shadow launch)
    int step_nodes, step_prims, min_active;   // persistent traversal: step bounds, refill threshold
    int node_min, prim_min;                   // persistent traversal: warp-vote thresholds of the inner loops (0 = plain step)
    int shade_ctas;         // k_shade_all variant: resident CTAs per SM its registers are held to (2, 3 or 4)
    bool walker;            // persistent walker kernels (else the flat per-thread traversal kernels)
    bool sort_shadow;       // option sort_rays >= 2: the NEE rays are binned too
    bool stream;            // streaming wavefront (kernels/stream.cu) instead of the slot pool (kernels/wavefront.cu)
    int grid_fill, grid_stream;   // its two grids
    cudaEvent_t mid_event;  // if set: recorded in the middle of the iteration (after the traversal / fill launch)
    const SmallScene* small; // host copy of a <= RTW_SMALL_MAX-leaf scene passed to the *_small kernels by value (else null)
};
cudaError_t wavefront_configure(const DScene& sc, LaunchCfg* cfg);
// One wavefront iteration: regen, reset counters, traverse, shade*, shadow.
cudaError_t wavefront_iteration(cudaStream_t s, const DScene& sc, const DWave& w, const LaunchCfg& cfg,
                                uint64_t* launches, cudaEvent_t* stage_events /* 5 or null */);
cudaError_t wavefront_init(cudaStream_t s, const DWave& w, uint64_t* launches);
// Streaming wavefront (kernels/stream.cu): fill + trace, shade + shadow + trace, advance.
cudaError_t stream_configure(const DScene& sc, LaunchCfg* cfg);
cudaError_t stream_init(cudaStream_t s, const DWave& w, uint64_t* launches);
cudaError_t stream_iteration(cudaStream_t s, const DScene& sc, const DWave& w, const LaunchCfg& cfg,
                             uint64_t* launches, cudaEvent_t* stage_events /* 5 or null */);
// Drain tail of a slot-pool frame in one launch pair (kernels/stream.cu): flush the slots that finished in the
// last iteration, then one thread per surviving path runs it to its end.  w.parity = parity of the iteration that
// is NOT run; w.trace_use = the sparse-mode list of survivors is valid (else the pool is scanned).
cudaError_t launch_tail(cudaStream_t s, const DScene& sc, const DWave& w, const LaunchCfg& cfg, uint64_t* launches);
cudaError_t launch_add_count(cudaStream_t s, int n_pixels, int count, float4* accum, uint64_t* launches);
cudaError_t launch_tonemap(cudaStream_t s, int n_pixels, int spp, const float4* accum, float4* out, uint64_t* launches);

// ---- GPU-side scene flattening / texture baking (kernels/bake.cu) ----
cudaError_t launch_invert_transforms(cudaStream_t s, const float* o2w, int n, rtw_transform* out);
cudaError_t launch_perlin_tables(cudaStream_t s, uint32_t seed, int first_table, int n_tables, float* draws /* 1533 * n_tables */, rtw_perlin* out);
cudaError_t launch_flip_rgba(cudaStream_t s, const uint8_t* src, int w, int h, int channels, uchar4* dst);

// ---- denoise stand-in (kernels/denoise.cu): `iterations` a-trous passes ping-ponging between a and b ----
cudaError_t launch_denoise(cudaStream_t s, int Nx, int Ny, int iterations, float sigma, float4* a, float4* b, float4** result,
                           uint64_t* launches);

// ---- test hooks --------------------------------------------------------------
cudaError_t launch_trace(cudaStream_t s, const DScene& sc, const LaunchCfg& cfg, const rtw_ray* rays, int n,
                         rtw_hit* hits, int32_t* occluded, uint32_t* work2 /* 2 device words */, uint64_t* launches);
cudaError_t launch_rng_stream(cudaStream_t s, uint32_t seed0, uint32_t seed1, int n, uint32_t* tea_out,
                              uint32_t* lcg_out, float* rnd_out, uint32_t* xs_out, float* randf_out);
cudaError_t launch_math_eval(cudaStream_t s, int fn, const float* x, const float* y, int n, float* out);
cudaError_t launch_camera_rays(cudaStream_t s, const DScene& sc, const DWave& w, int sample, rtw_ray* rays_out);

}  // namespace rtw
