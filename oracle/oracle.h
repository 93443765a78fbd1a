/* ORACLE — TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement of the reference's path-tracing launch
 * (reference RestOfLife/Director.cpp:982 → raygen/raygen.cu:123 and everything
 * below it).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.  The product
 * (raytracing-weekend_b200/) never includes, links or calls anything in oracle/.
 *
 * PARITY PINNED (round 2).  The reference ships no tests, golden vectors or fixtures for this path and its device
 * code needs optix.h — but its SOURCES compile for the host against the stand-in headers of oracle/ref_shim
 * (`make -C oracle ref` -> oracle/_ref/{ref_cli, libref.so, libref_dm.so, ref_rng}; nothing is copied).  The frames the
 * reference's own programs render there (scenes 0-4 through Director::init -> createScene -> renderFrame, plus three
 * hand-built scenes for the branches no built-in scene reaches) are committed as tests/golden/ref_frames.npz by
 * tests/golden/make_ref_golden.py, and tests/test_ref_pin.py asserts this oracle reproduces them BIT FOR BIT; tea /
 * xorshift32 / randf are pinned by lib/random.cuh compiled alone (tests/golden/rng_kat.json).  What the stand-in
 * decides where OptiX leaves a choice to the driver is listed at the top of oracle/ref_shim/optix.h.  Every function
 * below cites the reference file:line it restates.
 *
 * The scene description consumed here is the C-ABI POD of include/rtw.h.
 */
#ifndef ORACLE_H
#define ORACLE_H
#include <stdint.h>
#include "../include/rtw.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_scene orc_scene;

typedef struct orc_counters {
    uint64_t paths;
    uint64_t rays_closest;
    uint64_t rays_shadow;
    uint64_t nodes_visited;      /* BVH mode only */
    uint64_t prims_tested;
    uint64_t vertices_shaded;    /* closest-hit shader invocations */
    uint64_t shadow_enqueued;    /* == rays_shadow */
    /* split of nodes_visited / prims_tested (BVH mode), feeding the algorithmic
       bytes/ray of SURVEY.md §8d: closest-hit rays vs shadow rays, and how many of
       the tested primitives carry an instance transform (80 B instead of 32 B) */
    uint64_t nodes_closest, prims_closest, prims_xf_closest;
    uint64_t nodes_shadow, prims_shadow, prims_xf_shadow;
} orc_counters;

/* ---- RNG (lib/random.cuh:7-38; SDK cuda/random.h lcg/rnd) ---- */
uint32_t orc_tea(uint32_t rounds, uint32_t s0, uint32_t s1);
void orc_rng_stream(uint32_t seed0, uint32_t seed1, int n,
                    uint32_t* tea_out, uint32_t* lcg_out, float* rnd_out,
                    uint32_t* xs_out, float* randf_out);

/* ---- arithmetic contract (oracle/detmath.h): fn numbered like rtw_math_fn ---- */
void orc_math_eval(int fn, const float* x, const float* y, int n, float* out);

/* ---- scene ---- */
int  orc_scene_load(const rtw_scene_desc* desc, orc_scene** out);   /* deep copy */
void orc_scene_free(orc_scene* s);

/* ---- rays ---- */
void orc_camera_rays(const orc_scene* s, int Nx, int Ny, int sample, rtw_ray* out);
/* mode 0 = brute force over all primitives (ground truth), 1 = host-mirror LBVH */
void orc_trace_closest(const orc_scene* s, const rtw_ray* rays, int n, rtw_hit* hits,
                       int mode, orc_counters* counters);
void orc_trace_any(const orc_scene* s, const rtw_ray* rays, int n, int32_t* occluded,
                   int mode, orc_counters* counters);

/* ---- render: linear sums (r,g,b,count) per pixel, row 0 = bottom ---- */
int  orc_render(const orc_scene* s, int Nx, int Ny,
                int sample_first, int sample_stride, int sample_count,
                int max_depth, int n_threads, int mode,
                float* rgba_sum, orc_counters* counters);
/* rows [y0,y1) only (bounded CPU-baseline samples) */
int  orc_render_rows(const orc_scene* s, int Nx, int Ny, int y0, int y1,
                     int sample_first, int sample_stride, int sample_count,
                     int max_depth, int n_threads, int mode,
                     float* rgba_sum, orc_counters* counters);
/* out = (sqrt(sum/spp), 1)  (raygen/raygen.cu:144-158) */
void orc_tonemap(int n_pixels, int spp, const float* rgba_sum, float* rgba_out);
/* single sample radiance, for debugging and per-path tests */
void orc_sample_radiance(const orc_scene* s, int Nx, int Ny, int x, int y, int k,
                         int max_depth, int mode, float* rgb_out);

/* ---- host-mirror LBVH (same algorithm as the device builder) ---- */
int  orc_bvh_build(orc_scene* s);
/* subtrees of <= leaf_size leaves are tested as one primitive list (0 = the product's
   default: 16 for scenes of <= 16 primitives, else 4); affects the visit counters only */
void orc_bvh_set_leaf_size(orc_scene* s, int leaf_size);
int  orc_bvh_info(const orc_scene* s, rtw_bvh_info* info);
int  orc_bvh_download(const orc_scene* s, uint64_t* morton, int32_t* leaf_prim,
                      int32_t* left, int32_t* right, float* bounds);
/* per-primitive world AABBs as the builder sees them: 6 floats per prim */
void orc_prim_aabbs(const orc_scene* s, float* aabbs);

/* the product's denoise stand-in (csrc/kernels/denoise.cu), restated; NOT the reference's OptiX AI denoiser */
void orc_denoise(int Nx, int Ny, int iterations, float sigma, const float* rgba_in, float* rgba_out);

/* texture evaluation hook (tests) */
void orc_texture_eval(const orc_scene* s, int texture, float u, float v,
                      const float* p, float* rgb_out);

#ifdef __cplusplus
}
#endif
#endif
