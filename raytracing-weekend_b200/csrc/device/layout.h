// layout.h — HBM data layout of the scene, the LBVH and the wavefront state.
// (DESIGN.md "Data layout in HBM" describes every array and its bytes/unit.)
#pragma once
#include "vec.h"
#include "../../../include/rtw.h"

namespace rtw {

// ---- child reference of a BVH node ---------------------------------------
// bit 31: leaf RANGE — bits 0..24 = first leaf index (position in Morton order),
//         bits 25..29 = count - 1.  A subtree of the Karras hierarchy that spans at
//         most `leaf_size` (<= 32) consecutive leaves is referenced as one range and
//         its primitives are tested in a loop; count == 1 is the plain LBVH leaf.
// bit 30: node lives in the shared-memory "top" table (low bits = top index)
// else  : index into the global node array
#define RTW_REF_LEAF 0x80000000u
#define RTW_REF_TOP  0x40000000u
#define RTW_REF_MASK 0x3fffffffu
#define RTW_REF_NONE 0xffffffffu
#define RTW_LEAF_FIRST_MASK 0x01ffffffu
#define RTW_MAX_LEAF_SIZE 32
#define RTW_MAX_PRIMS (1 << 25)
RTW_HD uint32_t leaf_range_ref(int first, int count) {
    return RTW_REF_LEAF | ((uint32_t)(count - 1) << 25) | (uint32_t)first;
}
RTW_HD uint32_t leaf_range_first(uint32_t ref) { return ref & RTW_LEAF_FIRST_MASK; }
RTW_HD uint32_t leaf_range_count(uint32_t ref) { return ((ref >> 25) & 31u) + 1u; }

// ---- primitive record: 2 x float4 per leaf, stored in leaf (Morton) order ----
//   SPHERE         a = c.xyz r          b = -  -  -  meta
//   MOVING_SPHERE  a = c0.xyz r         b = c1.xyz   meta     (+ prim_ext: time0 time1)
//   RECT_X/Y/Z     a = a0 a1 b0 b1      b = k  -  -  meta
//   VOLUME_BOX     a = min.xyz density  b = max.xyz  meta
//   VOLUME_SPHERE  a = c.xyz r          b = density - - meta
// meta (bit pattern of b.w): type[0:3) flip[3] material-type[4:7) (transform+1)[7:32)
RTW_HD uint32_t meta_pack(int type, int flip, int mattype, int xform) {
    return (uint32_t)type | ((uint32_t)(flip & 1) << 3) | ((uint32_t)mattype << 4) | ((uint32_t)(xform + 1) << 7);
}
RTW_HD int meta_type(uint32_t m) { return (int)(m & 7u); }
RTW_HD int meta_flip(uint32_t m) { return (int)((m >> 3) & 1u); }
RTW_HD int meta_mattype(uint32_t m) { return (int)((m >> 4) & 7u); }
RTW_HD int meta_xform(uint32_t m) { return (int)(m >> 7) - 1; }  // -1 = identity

#define RTW_MAX_LIGHTS 8

struct DLight {            // LightDefinition (reference lib/raydata.cuh:32-49)
    float3 normal;
    float  area;
    float3 emission;
    float  pad;
};

struct DImage {
    const uchar4* texels;  // RGBA8, row 0 = v == 0
    int width, height;
};

// Read-only scene, passed to kernels by value (lives in the constant bank).
struct DScene {
    // LBVH
    const float4* nodes;        // 4 float4 (64 B) per internal node, see lbvh.h
    const float4* top_nodes;    // n_top nodes re-indexed breadth-first for shared-memory staging
    const float4* wide;         // 8 float4 (128 B) per internal node: the node collapsed to <= 4 children (lbvh.h make_wide_node);
                                // nullptr = walk the two-child nodes
    const float4* prims;        // 2 float4 (32 B) per leaf
    const float2* prim_ext;     // per leaf: (time0, time1) of a moving sphere
    const int*    leaf_prim;    // leaf → index into rtw_scene_desc::prims
    const int*    leaf_mat;     // leaf → material index
    const float4* xforms;       // 6 float4 per transform: world→object rows 0-2, object→world rows 0-2
    uint32_t root;              // child reference of the root
    int n_leaves, n_internal, n_top;
    int over_first, over_count; // oversize primitives (lbvh.h): leaves [over_first, over_first + over_count) are outside the
                                // hierarchy and tested by every ray before it walks; over_count == 0: none
    int has_media;              // scene contains constant media (ray seed needed in traversal)
    int has_motion;             // scene contains moving spheres (gather time needed by shadow rays)
    int has_images;             // scene contains image textures (sphere uv needed)
    uint32_t* overflow;         // one device word: traversal-stack pushes that found the stack full (must stay 0)
    // shading tables
    const float4* mats;         // 1 float4 per material: (type, texture, param, -) as bit patterns
    const float4* texs;         // 2 float4 per texture: (type, r, g, b) (odd, even, scale, table/image index)
    const float4* perlin_vec;   // 256 float4 per Perlin table
    const int*    perlin_perm;  // 768 ints per Perlin table: perm_x, perm_y, perm_z
    const DImage* images;
    // lights (reference SysParamter: lights, numLights, pdf; Director.cpp:523-549)
    int n_lights;
    int sky_light;
    int has_light_rect;
    int rect_axis;
    float rect_a0, rect_a1, rect_b0, rect_b1, rect_k;
    DLight lights[RTW_MAX_LIGHTS];
    // camera (shaders/sysparameter.h:44-56)
    float3 cam_origin, cam_u, cam_v, cam_ll, cam_horizontal, cam_vertical;
    float cam_lens_radius, cam_time0, cam_time1;
};

// A scene of <= RTW_SMALL_MAX leaves (the Cornell scenes of the reference: 13 and 8 primitives) is
// ONE leaf range, walked by every ray in the same order.  Its leaf records travel to the traversal
// kernels BY VALUE as a kernel parameter: the loop counter over the leaves is warp-uniform, so the
// records are read through the constant bank into uniform registers, the kind / transform dispatch
// is a uniform branch, and no lane spends a load or a decode on them.
#define RTW_SMALL_MAX 16
struct SmallLeaf {
    float4 a, b;                      // DScene::prims of the leaf
    float2 ext;                       // DScene::prim_ext
    int32_t prim_id;                  // DScene::leaf_prim
    int32_t leaf;                     // index of this record in the scene's leaf order (what the hit record stores)
};
// The records are ordered by (transform, kind) and cut into RUNS of equal (transform, kind): the ray is
// carried into a transform once per run and the inner loop of a run is ONE candidate test with no dispatch.
// The order of the tests does not matter: the closest hit is the lexicographic minimum of (t, primitive id).
struct SmallRun {
    int32_t first, count;             // records [first, first + count) of SmallScene::leaf
    int32_t kind;                     // RTW_PRIM_*
    int32_t xf;                       // transform id, -1 = none (always -1 for moving spheres)
};
// The STATIC part of a small scene: rects and (non-moving) spheres sit in a table of fixed shape
// [group][kind][slot], group = one transform (group 0 = untransformed).  The walk over it is fully
// unrolled, so every record field is an immediate constant-bank operand of the instruction that uses
// it (no index arithmetic, no load), and a run of an absent kind costs one uniform branch.
// Records that do not fit the shape (other kinds, more than RTW_SS_* of a kind) stay in the runs.
#define RTW_SS_GROUPS  3
#define RTW_SS_RECTS   4      // per group and axis
#define RTW_SS_SPHERES 2      // per group
struct SmallRect {
    float4 a;                         // bounds (lo, hi, lo, hi) of the two in-plane axes
    float k;                          // plane offset
    int32_t prim_id;
    int32_t leaf;                     // index in the scene's leaf order
    int32_t pad;
};
struct SmallSphere {
    float4 c;                         // centre, radius
    int32_t prim_id, leaf, pad[2];
};
struct SmallGroup {
    float4 w2o[3];                    // world→object rows (unused for group 0)
    SmallRect rect[3][RTW_SS_RECTS];  // [axis: x, y, z]
    SmallSphere sphere[RTW_SS_SPHERES];
    int32_t n_rect[3];
    int32_t n_sphere;
};
struct SmallScene {
    SmallGroup group[RTW_SS_GROUPS];
    SmallLeaf leaf[RTW_SMALL_MAX];    // the records left to the runs
    float4 w2o[3 * RTW_SMALL_MAX];    // world→object rows of transform j (first half of DScene::xforms), for the runs
    SmallRun run[RTW_SMALL_MAX];
    int32_t n;                        // leaves of the scene (static + runs)
    int32_t n_runs;
    int32_t n_groups;                 // groups in use (0 = nothing static)
    int32_t pad;
};

// ---- wavefront state: structure of arrays over P path slots ---------------
// bytes per slot: ray 32 + throughput/seed 16 + misc 8 + hit 8 (one 64-byte record) + radiance/pixel 16 = 80
#define RTW_ST_EMPTY 0u
#define RTW_ST_ALIVE 1u
#define RTW_ST_DEAD  2u     // finished: radiance waits to be flushed by k_logic

// shading queue classes ("one BSDF per warp")
#define RTW_Q_LAMBERTIAN 0
#define RTW_Q_METAL      1
#define RTW_Q_DIELECTRIC 2
#define RTW_Q_ISOTROPIC  3
#define RTW_Q_TERMINAL   4   // miss, diffuse light, normal material: path ends
#define RTW_Q_COUNT      5   // shading classes
#define RTW_Q_REGEN      5   // slots whose path finished this iteration: flushed and refilled by k_regen
#define RTW_Q_TRACE      6   // sparse mode only: the slots that hold a ray to trace in the next traversal
#define RTW_Q_SLOTS      7   // queues allocated (RTW_Q_SLOTS * P slot indices)

// counters[] layout (uint32 unless noted)
#define RTW_C_QUEUE0     0                 // RTW_Q_SLOTS queue lengths (5 shading classes + regen)
#define RTW_C_SHADOW     6                 // shadow queue length
#define RTW_C_ALIVE      7                 // closest-hit rays traced by the last iteration (0 = frame done)
#define RTW_C_NEXT_LO    8                 // uint64 next sample ordinal (two words)
#define RTW_C_RAYS_LO    10                // uint64 closest-hit rays traced
#define RTW_C_SHRAYS_LO  12                // uint64 shadow rays traced
#define RTW_C_PATHS_LO   14                // uint64 paths flushed
#define RTW_C_WORK_CLOSEST 16              // next pool slot to fetch (persistent traversal warps)
#define RTW_C_WORK_SHADOW  17              // next shadow-queue entry to fetch
#define RTW_C_TRACE0     18                // sparse mode: length of the trace list of even / odd iterations (two words)
// streaming wavefront (device/stream.h): queue lengths [buffer][class], the paths the next fill starts,
// the closest-hit rays the shading kernel traced this iteration
#define RTW_C_SQ0        20                // 2 * 4 words
#define RTW_C_NEW        28
#define RTW_C_TRACED     29
#define RTW_C_TICKET     30                // CTAs of k_regen that have finished (the last one retires the iteration's counters)
#define RTW_C_UNOCC_LO   32                // uint64 NEE contributions added (unoccluded shadow rays)
#define RTW_C_WORDS      36

// Per-slot state is read and written through queue indices (a warp of the shading kernels holds
// 32 arbitrary slots), so the records that are always used together are laid out as ONE 32-byte
// sector each: a slot costs exactly the bytes it uses, whatever its neighbours do, and a record
// moves with one 256-bit access (LDG.E.ENL2.256 / STG.E.ENL2.256).
struct __align__(32) RayRec {
    float4 o;           // origin.xyz, ray time   (o.w == RTW_RAY_NONE bits: the slot holds no live ray)
    float4 d;           // direction.xyz, gather time
};
struct __align__(32) PathRec {
    float4 thr;         // throughput.xyz, path seed bits
    uint2 misc;         // (depth | state << 8, ray-time seed)
    uint2 hit;          // (t bits, leaf index or RTW_REF_NONE) of the slot's current ray: written by the closest-hit stage, read by shading
};
struct __align__(64) SlotRec {   // one 64-byte line per slot: L2 fetches DRAM in 64-byte pairs of sectors (the
    RayRec ray;                  // default cudaLimitMaxL2FetchGranularity), so a shading warp that gathers ray +
    PathRec st;                  // path state through queue indices pays for exactly the bytes it uses
};
#define RTW_CAMERA_BINS (6 * 65536)   // option sort_rays: direction bins of the camera rays (cube map, 256 x 256 per face)
#define RTW_RAY_NONE   0xffffffffu   // RayRec.o.w of a slot without a live ray (a ray time is a number in [0, 1))
#define RTW_PIXEL_NONE 0xffffffffu   // rad.w of a slot that holds nothing to flush

struct DWave {
    SlotRec* rec;       // current ray + path state of the slot
    float4* rad;        // radiance sum .xyz, pixel index bits (touched only when light is added)
    uint32_t* queue;    // RTW_Q_SLOTS * P slot indices
    float4* sh_o;       // shadow queue: origin.xyz, tmax
    float4* sh_d;       //               direction.xyz, slot bits
    float4* sh_c;       //               contribution.xyz, path seed bits
    float4* sq;         // streaming wavefront: 2 buffers x 4 classes x 5 float4 arrays of P entries (device/stream.h)
    uint32_t* counters; // RTW_C_WORDS words
    // Option sort_rays (LBVH walkers): a stage takes its rays bin by bin — bin = (origin cell, direction bin) — instead
    // of in slot / queue order.  One counting sort per stage and iteration (kernels/wavefront.cu k_ray_bin, k_bin_scan,
    // k_ray_place): bins[key] counts, then holds the first position of the bin; bin_rank[i] = (key, rank inside the bin)
    // of item i; order[position] = item.  bins[n_bins] = number of items placed (empty slots are left out).
    uint32_t* bins;             // n_bins + 1 + ceil(n_bins / 1024) words (counts / first positions, the total, sums of 1024-bin chunks); nullptr = option off
    int n_bins;                 // 2^(3 * (key_obits + key_dbits)) regular bins + (key_camera ? the camera rays' direction bins : 0)
    int key_camera;             // camera rays (depth 0) are binned by direction alone, in bins of their own (kernels/wavefront.cu camera_key)
    uint2* bin_rank;            // P entries
    uint32_t* order;            // P entries
    int use_order;              // this launch walks `order` (set by the iteration for the launch that has one)
    int key_obits, key_dbits;   // key = Morton code of the origin cell (obits per axis) << 3 * dbits | direction bin (dbits per axis)
    float3 key_lo, key_scale;   // origin -> cell: (o - key_lo) * key_scale in [0, 2^obits)
    float4* accum;      // Nx*Ny linear radiance sums (rgb) + sample count (w)
    int P;
    int Nx, Ny;
    int max_depth;
    int sample_first, sample_stride;
    unsigned long long total;   // pixel_count*sample_count sample ordinals to issue
    // Image tiles (multi-GPU row sharding, api/rtw_multi.cu): the call renders the pixels
    // [pixel_first, pixel_first + pixel_count) of the Nx x Ny image — whole rows — into the full-size accumulator.
    uint32_t pixel_first, pixel_count;
    // Sparse mode (the drain tail of a frame, when most slots are empty; set per iteration by the host):
    // the slots with a ray to trace are listed instead of found by scanning the pool.
    int parity;                 // iteration index & 1
    int trace_fill;             // shading and regeneration list the slots the NEXT / this traversal must trace
    int trace_use;              // this iteration's traversal walks the list (filled since the previous iteration)
    int first_iter;             // first iteration of a frame: every slot is free and holds nothing (k_regen takes slot i, reads nothing)
};

}  // namespace rtw
