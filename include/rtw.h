/*
 * rtw.h — C-ABI of the B200-native wavefront path tracer.
 *
 * This header is the drop-in boundary of the hot path.  The
 * reference has no FFI of its own; the narrowest seam is the `Director` class
 * (reference RestOfLife/Director.h:118-126) driven from RestOfLife/main.cpp:145-164.
 * Every entry point below names the reference interface it replaces.
 *
 * Conventions
 *   - plain C: pointers + sizes, no C++/torch types, no exceptions, no exit().
 *   - every function returns an int status (RTW_OK == 0); rtw_last_error()
 *     returns a human readable string for the last failure on that context.
 *   - host arrays passed in are borrowed for the duration of the call only.
 *   - a context is bound to ONE CUDA device and is not thread-safe.
 *   - there is NO CPU fallback: without a usable CUDA device rtw_create fails
 *     with RTW_ERR_NO_DEVICE.
 *
 * All arithmetic is fp32, RNG state is uint32.
 */
#ifndef RTW_H
#define RTW_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTW_ABI_VERSION 2

/* ---- status codes ------------------------------------------------------ */
#define RTW_OK             0
#define RTW_ERR_INVALID    1   /* bad argument                               */
#define RTW_ERR_CUDA       2   /* a CUDA runtime call failed                 */
#define RTW_ERR_NO_DEVICE  3   /* no usable CUDA device (no CPU fallback)    */
#define RTW_ERR_STATE      4   /* call order violated (e.g. render before build) */
#define RTW_ERR_IO         5   /* file could not be read / written           */

typedef struct rtw_ctx rtw_ctx;     /* opaque: one CUDA device, one stream set */

/* ---- enumerations ------------------------------------------------------ */

/* Primitive kinds == the reference's intersection programs
 * (geometry/sphere.cu:48, geometry/movingSphere.cu:61, shaders/aarect{x,y,z}.cu:8,
 *  geometry/volumeBox.cu:55, geometry/volumeSphere.cu:67). */
enum rtw_prim_type {
    RTW_PRIM_SPHERE        = 0,
    RTW_PRIM_MOVING_SPHERE = 1,
    RTW_PRIM_RECT_X        = 2,
    RTW_PRIM_RECT_Y        = 3,
    RTW_PRIM_RECT_Z        = 4,
    RTW_PRIM_VOLUME_BOX    = 5,
    RTW_PRIM_VOLUME_SPHERE = 6,
    RTW_PRIM_TYPE_COUNT    = 7
};

/* Material kinds, numbered like MaterialType (lib/raydata.cuh:23-30). */
enum rtw_material_type {
    RTW_MAT_LAMBERTIAN    = 0,
    RTW_MAT_DIFFUSE_LIGHT = 1,
    RTW_MAT_METAL         = 2,
    RTW_MAT_DIELECTRIC    = 3,
    RTW_MAT_ISOTROPIC     = 4,
    RTW_MAT_NORMAL        = 5,
    RTW_MAT_TYPE_COUNT    = 6
};

/* Texture kinds, numbered like TexCallFunction (shaders/FunctionIdx.h:8-15). */
enum rtw_texture_type {
    RTW_TEX_CHECKER  = 0,
    RTW_TEX_CONSTANT = 1,
    RTW_TEX_IMAGE    = 2,
    RTW_TEX_NOISE    = 3,
    RTW_TEX_NULL     = 4,
    RTW_TEX_TYPE_COUNT = 5
};

/* ---- scene description (POD, replaces HitGroupData / MaterialParams /
 *      textureParam / LightDefinition / SysParamter camera fields) --------- */

/* One primitive (reference: one OptixInstance + one HitGroupData record,
 * lib/raydata.cuh:95-115).  Layout of v[] by type:
 *   SPHERE         v = cx cy cz r
 *   MOVING_SPHERE  v = c0x c0y c0z r c1x c1y c1z time0 time1
 *   RECT_X/Y/Z     v = a0 a1 b0 b1 k            (hitRectData, raydata.cuh:76-84)
 *   VOLUME_BOX     v = minx miny minz maxx maxy maxz density
 *   VOLUME_SPHERE  v = cx cy cz r density
 */
typedef struct rtw_prim {
    int32_t type;       /* rtw_prim_type                                       */
    int32_t material;   /* index into materials (reference: instanceId → MatParams) */
    int32_t transform;  /* index into transforms, -1 = identity                 */
    int32_t flip;       /* rect only: normal = -axis (hitRectData.flip)        */
    float   v[12];
} rtw_prim;             /* 64 bytes */

/* Instance transform: row-major 3x4 object→world (OptixInstance::transform) and
 * its inverse world→object (computed by OptiX internally in the reference; by
 * the host scene builder here). */
typedef struct rtw_transform {
    float o2w[12];
    float w2o[12];
} rtw_transform;

/* MaterialParams (shaders/sysparameter.h:5-15). */
typedef struct rtw_material {
    int32_t type;       /* rtw_material_type */
    int32_t texture;    /* index into textures, -1 = none (dielectric, normal) */
    float   param;      /* metal: fuzz (clamped to 1, material/ioMetalMaterial.h:28-32); dielectric: eta */
    int32_t pad;
} rtw_material;

/* textureParam (lib/raydata.cuh:127-138). */
typedef struct rtw_texture {
    int32_t type;       /* rtw_texture_type */
    float   color[3];   /* constant */
    int32_t odd, even;  /* checker: indices into textures (Q18: dispatched correctly) */
    float   scale;      /* noise */
    int32_t perlin;     /* noise: index into perlin tables */
    int32_t image;      /* image: index into images */
    int32_t pad[3];
} rtw_texture;          /* 48 bytes */

/* Perlin tables (texture/ioTexture.h:181-216). */
typedef struct rtw_perlin {
    float   ranvec[256][3];
    int32_t perm_x[256];
    int32_t perm_y[256];
    int32_t perm_z[256];
} rtw_perlin;

/* RGBA8 image, row 0 = v == 0 (already flipped like texture/ioTexture.h:247-266). */
typedef struct rtw_image {
    int32_t width, height;
    const uint8_t* rgba;    /* width*height*4 bytes */
} rtw_image;

/* LightDefinition (lib/raydata.cuh:32-49). */
typedef struct rtw_light {
    float position[3];
    float vecU[3];
    float vecV[3];
    float normal[3];
    float area;
    float emission[3];
} rtw_light;

/* The single sampling rectangle of pdf.p1 (scene/ioScene.h:103-148, Q26). */
typedef struct rtw_light_rect {
    int32_t axis;           /* 0 = X, 1 = Y, 2 = Z (rect_{x,y,z}_generate) */
    float   a0, a1, b0, b1, k;
} rtw_light_rect;

/* Camera fields of SysParamter (shaders/sysparameter.h:44-56). */
typedef struct rtw_camera {
    float origin[3];
    float u[3], v[3], w[3];
    float lower_left[3];
    float horizontal[3];
    float vertical[3];
    float lens_radius;      /* 0 in every reference scene (Q3) */
    float time0, time1;
} rtw_camera;

typedef struct rtw_scene_desc {
    int32_t n_prims;       const rtw_prim*      prims;
    int32_t n_transforms;  const rtw_transform* transforms;
    int32_t n_materials;   const rtw_material*  materials;
    int32_t n_textures;    const rtw_texture*   textures;
    int32_t n_perlin;      const rtw_perlin*    perlin;
    int32_t n_images;      const rtw_image*     images;
    int32_t n_lights;      const rtw_light*     lights;
    int32_t has_light_rect; rtw_light_rect      light_rect;
    rtw_camera camera;
    int32_t sky_light;      /* Director.cpp:523: true iff the scene has no lights */
    int32_t reserved;
} rtw_scene_desc;

/* ---- test-hook PODs ---------------------------------------------------- */
typedef struct rtw_ray {
    float origin[3];
    float tmin;
    float dir[3];
    float tmax;
    float time;         /* ray time (motion transform key interpolation)      */
    float gather_time;  /* PerRayData::gatherTime (moving sphere centre)      */
    uint32_t seed;      /* path seed at ray start (constant-medium free flight, Q23) */
    uint32_t pad;
} rtw_ray;              /* 48 bytes */

typedef struct rtw_hit {
    float   t;          /* hit distance, undefined when prim < 0 */
    int32_t prim;       /* index into rtw_scene_desc::prims, -1 = miss */
} rtw_hit;

typedef struct rtw_stats {
    uint64_t paths;             /* camera samples completed by the last render */
    uint64_t rays_closest;      /* closest-hit traversals                       */
    uint64_t rays_shadow;       /* any-hit (shadow) traversals                  */
    uint64_t kernel_launches;   /* kernels launched by the last render          */
    uint64_t iterations;        /* wavefront iterations of the last render, summed over its waves */
    double   render_ms;         /* device time of the last render (CUDA events) */
    double   build_ms;          /* device time of the last rtw_build_accel      */
    /* option "profile": CUDA events around the stages of sampled iterations of wave 0, scaled to all
     * iterations of all waves — i.e. summed kernel time; with more than one wave the launches of different
     * waves overlap on the device, so the four figures add up to more than render_ms */
    double   traverse_ms;       /* device time inside k_traverse_closest, if profiling was enabled, else 0 */
    double   shadow_ms;         /* device time inside k_traverse_shadow, ditto  */
    double   shade_ms;          /* device time inside the shading kernels, ditto */
    double   logic_ms;          /* device time inside k_logic, ditto            */
    uint64_t waves;             /* concurrent pools the last render was dealt to */
    uint64_t pool_paths;        /* path slots per pool                          */
    uint64_t nee_unoccluded;    /* NEE contributions added (shadow rays that reached the light); 0 where the variant does not count them */
    uint64_t stack_overflows;   /* traversal-stack pushes that found the stack full (must be 0: the render returns RTW_ERR_STATE otherwise) */
} rtw_stats;

typedef struct rtw_bvh_info {
    int32_t n_leaves;
    int32_t n_internal;
    int32_t root;
    int32_t n_top;              /* internal nodes staged in shared memory       */
    float   scene_min[3];
    float   scene_max[3];
} rtw_bvh_info;

/* ---- context lifetime (replaces Director::init / destroy,
 *      Director.cpp:33-104, Director.h:120,126) --------------------------- */
int  rtw_abi_version(void);
int  rtw_device_count(int* count);
int  rtw_create(int device, rtw_ctx** out);
int  rtw_destroy(rtw_ctx* ctx);
const char* rtw_last_error(const rtw_ctx* ctx);   /* ctx may be NULL: last create error */

/* ---- scene + acceleration structure (replaces Director::createScene →
 *      createSBT + initLaunchParams, Director.cpp:483-553,628-885, and every
 *      optixAccelBuild call site) -------------------------- */
int  rtw_upload_scene(rtw_ctx* ctx, const rtw_scene_desc* scene);
int  rtw_build_accel(rtw_ctx* ctx);

/* ---- render (replaces Director::renderFrame's optixLaunch,
 *      Director.cpp:971-986 → raygen/raygen.cu:123) ----------------------- */

/* Synchronous.  rgba_out_host: caller-owned Nx*Ny*4 floats, row 0 = bottom row,
 * sqrt-gamma encoded, alpha 1 (raygen/raygen.cu:151-158).  spp == 1 reproduces
 * the reference launch (Q1); spp > 1 averages samples k = 0..spp-1 whose seeds
 * are tea<64>(pixel, k).  Option "seed_base" = b shifts the range to k = b..b+spp-1
 * (SURVEY.md §8b seed_base: disjoint ranges give independent estimates). */
int  rtw_render(rtw_ctx* ctx, int Nx, int Ny, int spp, int max_depth,
                float* rgba_out_host);

/* Multi-GPU building blocks.  accum_dev is a DEVICE buffer of Nx*Ny float4
 * holding linear (pre-gamma) radiance sums; it is NOT cleared by the call.
 * Samples rendered: k = sample_first + i*sample_stride, i = 0..sample_count-1.
 * Rank g of G uses sample_first = g, sample_stride = G.
 * Synchronous with respect to the host. */
int  rtw_render_accum(rtw_ctx* ctx, int Nx, int Ny,
                      int sample_first, int sample_stride, int sample_count,
                      int max_depth, float* accum_dev);
/* The same for an image TILE: only the rows [row_first, row_first + row_count) are rendered (all the
 * samples named by sample_first/stride/count), into the FULL-SIZE accumulator accum_dev (Nx*Ny float4).
 * A pixel's samples do not depend on the tile it is rendered in. */
int  rtw_render_accum_rows(rtw_ctx* ctx, int Nx, int Ny, int row_first, int row_count,
                           int sample_first, int sample_stride, int sample_count,
                           int max_depth, float* accum_dev);
/* out_dev[i] = (sqrt(accum.rgb / spp_total), 1)  (raygen/raygen.cu:144-158) */
int  rtw_tonemap(rtw_ctx* ctx, int Nx, int Ny, int spp_total,
                 const float* accum_dev, float* rgba_out_dev);
/* Copy the linear sums of the last rtw_render (for relMSE, SURVEY.md §8b "Output"). */
int  rtw_get_linear(rtw_ctx* ctx, float* rgba_linear_mean_host, int n_floats);

/* Tunables: "pool_paths" (wavefront pool size), "top_nodes" (BVH nodes staged in
 * shared memory), "profile" (per-stage event timing 0/1), "graph" (CUDA graph 0/1). */
int  rtw_set_option(rtw_ctx* ctx, const char* name, int64_t value);
int  rtw_get_stats(rtw_ctx* ctx, rtw_stats* out);

/* ---- the step between scene description and launch, on the device ------------
 * The reference inverts instance transforms inside optixAccelBuild and bakes its textures on the host, serially
 * (Director.cpp:498-519 -> texture/ioTexture.h:181-216 Perlin tables, :228-266 image decode / flip / upload).
 * These entry points do the same three steps as CUDA kernels; their outputs go straight into an rtw_scene_desc and
 * are BIT-EQUAL to what the host scene builder (rtw_scene_create) computes.  Host buffers in and out.
 *   rtw_bake_transforms  o2w_rows: n x 12 floats (row-major 3x4 object->world); out[i] = {o2w, w2o}
 *   rtw_bake_perlin      tables first_table .. first_table + n_tables - 1 of the stream of std::mt19937(mt_seed)
 *                        + uniform_real_distribution<float>(0,1) (the reference seeds 0; a table takes 1533 draws)
 *   rtw_bake_image       pixels: h rows top-down, `channels` (3 or 4) bytes per pixel; rgba_out: w*h*4 bytes,
 *                        row 0 = bottom row of the picture (v == 0), alpha 255 for 3-channel input */
int  rtw_bake_transforms(rtw_ctx* ctx, const float* o2w_rows, int n, rtw_transform* out);
int  rtw_bake_perlin(rtw_ctx* ctx, uint32_t mt_seed, int first_table, int n_tables, rtw_perlin* out);
int  rtw_bake_image(rtw_ctx* ctx, const uint8_t* pixels, int width, int height, int channels, uint8_t* rgba_out);

/* ---- the step after the path: a denoise STAND-IN (replaces Director.cpp:887-949,988-1000) ----------
 * The reference passes its frame through the OptiX AI denoiser, a closed network inside the driver that
 * cannot be matched; this entry point is a documented, DIFFERENT filter (edge-avoiding a-trous wavelet,
 * `iterations` passes 1..8 with tap spacing 2^i, colour weight 1 / (1 + d^2 / sigma^2)) on the same
 * gamma-encoded float4 frame, excluded from parity with the reference and never applied by rtw_render.
 * Host buffers in and out (Nx*Ny*4 floats each, may alias). */
int  rtw_denoise(rtw_ctx* ctx, int Nx, int Ny, int iterations, float sigma,
                 const float* rgba_in_host, float* rgba_out_host);

/* ---- the same filter without any allocation -----------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * rtw_denoise_static computes what rtw_denoise computes, bit for bit, on DEVICE buffers the caller owns
 * (Nx*Ny float4 each, on the context's device; they must NOT overlap: the input is read-only) and keeps its
 * ping-pong buffer in a file-scope `static __device__` array inside the library
 * (csrc/kernels/denoise.cu: RTW_DENOISE_STATIC_MAX_PIXELS float4 = 8 MiB, held by every GPU that loads the
 * library until the process exits).  It calls no cudaMalloc / cudaFree / allocating routine and does not
 * synchronise: the `iterations` passes are enqueued on `cuda_stream` (a cudaStream_t; NULL = the context's
 * own stream) and the call returns.  Nx*Ny > RTW_DENOISE_STATIC_MAX_PIXELS is refused with RTW_ERR_INVALID
 * (rtw_last_error names the limit).  The first call on a device looks the array's address up
 * (cudaGetSymbolAddress); make that call outside a stream capture.  All calls on one device share the
 * array: calls that overlap on the GPU (different streams or host threads) RACE — serialise them.  The
 * array keeps the last call's contents; nothing in it has to start at zero (every pass overwrites all
 * Nx*Ny pixels of its destination before they are read). */
#define RTW_DENOISE_STATIC_MAX_PIXELS (1 << 19)
int  rtw_denoise_static(rtw_ctx* ctx, int Nx, int Ny, int iterations, float sigma,
                        const float* rgba_in_dev, float* rgba_out_dev, void* cuda_stream);

/* ---- several GPUs of one box (SURVEY.md §8b `rtw_create(const int* devices, int n, ...)`, §8e) ----
 * One context, one worker thread and one stream set per listed GPU; scene and LBVH are replicated.
 * rtw_multi_render shards a frame by SAMPLE (GPU g renders samples k = g (mod G); one
 * ncclReduce(sum, float, root 0) of the fp32 accumulators per frame) or by image TILE (GPU g renders the
 * rows [g*Ny/G, (g+1)*Ny/G); ncclSend/ncclRecv of the row blocks to GPU 0).  Option "shard": 0 = auto
 * (tiles for frames below 2^23 paths, samples otherwise), 1 = samples, 2 = tiles; every other option is
 * forwarded to each context (rtw_set_option).  The image for any G equals the 1-GPU image up to the
 * order of the fp32 additions (bit for bit in tile mode at 1 spp).  The reference is single-GPU
 * (Director.cpp:971-986): these entry points replace the same Director calls as their 1-GPU twins.
 * NCCL is looked up at run time; without it (or with one GPU listed twice) peer copies are used. */
typedef struct rtw_multi rtw_multi;
int  rtw_create_multi(const int* devices, int n, rtw_multi** out);
int  rtw_multi_destroy(rtw_multi* m);
const char* rtw_multi_last_error(const rtw_multi* m);
int  rtw_multi_device_count(const rtw_multi* m);
int  rtw_multi_uses_nccl(const rtw_multi* m);
int  rtw_multi_set_option(rtw_multi* m, const char* name, int64_t value);
int  rtw_multi_upload_scene(rtw_multi* m, const rtw_scene_desc* scene);
int  rtw_multi_build_accel(rtw_multi* m);
int  rtw_multi_render(rtw_multi* m, int Nx, int Ny, int spp, int max_depth, float* rgba_out_host);
/* linear means of the last frame (w = sample count), as rtw_get_linear */
int  rtw_multi_get_linear(rtw_multi* m, float* rgba_linear_mean_host, int n_floats, int spp);
/* total: counters summed over the GPUs, render_ms = the slowest GPU; per_device: one rtw_stats per GPU
 * (may be NULL); exchange_ms: device time of the per-frame reduce / gather; shard: 1 samples, 2 tiles */
int  rtw_multi_get_stats(rtw_multi* m, rtw_stats* total, rtw_stats* per_device, double* exchange_ms, int* shard);

/* ---- tier-1 test hooks ---------------------------- */
/* BVH closest hit / any hit on a caller-supplied ray set (host arrays). */
int  rtw_trace_closest(rtw_ctx* ctx, const rtw_ray* rays, int n, rtw_hit* hits);
int  rtw_trace_any(rtw_ctx* ctx, const rtw_ray* rays, int n, int32_t* occluded);
/* RNG streams computed on the device: s = tea<64>(seed0, seed1); then n steps
 * of lcg → lcg_out[i], rnd_out[i] (SDK cuda/random.h as restated in SURVEY §8c);
 * and n steps of xorshift32/randf from the same start → xs_out[i], randf_out[i]
 * (lib/random.cuh:22-38).  Any output pointer may be NULL. */
int  rtw_rng_stream(rtw_ctx* ctx, uint32_t seed0, uint32_t seed1, int n,
                    uint32_t* tea_out, uint32_t* lcg_out, float* rnd_out,
                    uint32_t* xs_out, float* randf_out);
/* The path's transcendental functions as the device computes them (the arithmetic
 * contract of DESIGN.md; they replace CUDA libm's sinf/cosf/logf/powf(x,5)/atan2f/asinf
 * at the call sites lib/sampling.cuh:16-19,55-56, geometry/volumeBox.cu:79,
 * material/dielectricMaterial.cu:25, geometry/sphere.cu:27-28).  out[i] = fn(x[i], y[i]);
 * y is only read by RTW_FN_ATAN2 (out = atan2(x = ordinate, y = abscissa)) and may be NULL otherwise. */
enum rtw_math_fn { RTW_FN_SIN = 0, RTW_FN_COS = 1, RTW_FN_LOG = 2, RTW_FN_POW5 = 3, RTW_FN_ATAN2 = 4, RTW_FN_ASIN = 5,
                   RTW_FN_COUNT = 6 };
int  rtw_math_eval(rtw_ctx* ctx, int fn, const float* x, const float* y, int n, float* out);
/* Sample-0 camera rays as the device generates them (raygen/raygen.cu:129-135,
 * shaders/camera.cu:11-19); rays_out has Nx*Ny entries. */
int  rtw_camera_rays(rtw_ctx* ctx, int Nx, int Ny, int sample, rtw_ray* rays_out);
/* LBVH topology for parity against the host mirror.  Arrays are caller-owned:
 * morton[n_leaves] (sorted 63-bit codes), leaf_prim[n_leaves] (prim id of leaf i),
 * left/right[n_internal] (child refs: >=0 internal index, <0 → leaf ~ref),
 * bounds[n_internal*12] (lmin.xyz lmax.xyz rmin.xyz rmax.xyz).  Any may be NULL. */
int  rtw_bvh_info_get(rtw_ctx* ctx, rtw_bvh_info* info);
int  rtw_bvh_download(rtw_ctx* ctx, uint64_t* morton, int32_t* leaf_prim,
                      int32_t* left, int32_t* right, float* bounds);
/* Device radix sort hook (the LBVH builder's sort), host arrays in/out. */
int  rtw_sort_pairs(rtw_ctx* ctx, uint64_t* keys, uint32_t* values, int n);

/* ---- host-side scene flattening (replaces ioScene::init and the scene
 *      functions, scene/ioScene.h:52-101,180-982; Director.cpp:483-553) ---- */
typedef struct rtw_scene rtw_scene;
/* scene_id 0..4 = the reference's scenes; 5 = the synthetic stress scene C5
 * (n_stress_prims primitives, 0 → 2^20).  asset_dir holds
 * earthmap.ppm for scenes 2 and 4 (NULL → "assets"). */
int  rtw_scene_create(int scene_id, int Nx, int Ny, const char* asset_dir,
                      int n_stress_prims, rtw_scene** out);
const rtw_scene_desc* rtw_scene_get_desc(const rtw_scene* s);
const char* rtw_scene_description(const rtw_scene* s);
void rtw_scene_free(rtw_scene* s);

/* ---- scene wire format (the step before the path: the reference's scenes are C++
 *      functions, scene/ioScene.h:180-982; here a flattened scene is a file) ------
 * rtw_scene_save writes `desc` (+ an optional description text) as one little-endian
 * file: magic "RTWS", version, header, the PODs above array by array, FNV-1a
 * checksum.  rtw_scene_load reads it back into an rtw_scene, to be freed with
 * rtw_scene_free, after validating sizes, checksum and every index a kernel follows.
 * RTW_ERR_IO: cannot open / short file; RTW_ERR_INVALID: not a scene file, wrong
 * version, checksum mismatch, index out of range.  rtw_scene_io_error(): the reason. */
int  rtw_scene_save(const rtw_scene_desc* desc, const char* description, const char* path);
int  rtw_scene_load(const char* path, rtw_scene** out);
const char* rtw_scene_io_error(void);

/* ---- output (replaces Director::printPPM, Director.cpp:1010-1031) ------ */
/* P3 text, top row first, int(255.99f*clamp(c,0,1)); path NULL or "-" → stdout. */
int  rtw_write_ppm(const char* path, int Nx, int Ny, const float* rgba);

#ifdef __cplusplus
}
#endif
#endif /* RTW_H */
