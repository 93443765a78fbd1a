// ORACLE — TEST INFRASTRUCTURE ONLY (oracle/ref_shim).
//
// optix.h — a CPU stand-in for the subset of the OptiX 8.0.0 host and device API that the
// reference uses (grep over /root/reference/RestOfLife: 35 device entry points, 30 host
// calls), so that g++ compiles the reference's own main.cpp, Director.cpp, scene/, geometry/,
// material/, texture/, pdf/, shaders/, raygen/, miss/ sources WHERE THEY LIE and runs them on
// the host: oracle/_ref/ref_cli and oracle/_ref/libref*.so (recipe: oracle/Makefile `ref`).
// Nothing here is derived from NVIDIA's headers beyond the public names, struct fields and
// documented semantics the reference's call sites need; the implementation is
// oracle/ref_shim/ref_runtime.cpp.
//
// Runtime semantics supplied by the stand-in (OptiX leaves them to the driver; the oracle
// states the same rules, oracle/oracle.cpp "Intersection programs"):
//   - traversal is a loop over the instances of the IAS, last to first, so that among equal
//     hit distances the lowest instance index wins; no AABB culling (the AABBs handed to
//     optixAccelBuild are kept for inspection only);
//   - optixReportIntersection accepts iff tmin <= t <= tmax and then shrinks tmax;
//   - an instance whose 3x4 transform is exactly the identity leaves the ray untouched;
//     otherwise object ray = W2O * world ray, W2O = the inverse formed in double by cofactors
//     and rounded once; normals go to world space with the transpose of W2O;
//   - a matrix motion transform with identity rotation (the only kind the reference builds,
//     geometry/ioMovingSphere.h:162-182) translates by (1-f)*key0 + f*key1, f = ray time
//     clamped to [timeBegin, timeEnd];
//   - tex2D<float4> = normalised coordinates, clamp, fp32 bilinear on RGBA8 (x = u*W - 0.5);
//   - the denoiser is the identity (the OptiX AI denoiser is a closed network: out of scope).
#pragma once
#include "vector_types.h"
#include "texture_types.h"
#include "cuda_runtime.h"

// ---------------------------------------------------------------------------------------
// host-side types
// ---------------------------------------------------------------------------------------
typedef unsigned long long CUdeviceptr;
typedef unsigned long long OptixTraversableHandle;
typedef unsigned int OptixVisibilityMask;
typedef struct CUctx_st* CUcontext;

typedef enum OptixResult { OPTIX_SUCCESS = 0, OPTIX_ERROR_INVALID_VALUE = 7001 } OptixResult;

#define OPTIX_SBT_RECORD_HEADER_SIZE ((size_t)32)
#define OPTIX_SBT_RECORD_ALIGNMENT 16ull

namespace refshim { struct Context; struct Module; struct ProgramGroup; struct Pipeline; struct Denoiser; }
typedef refshim::Context* OptixDeviceContext;
typedef refshim::Module* OptixModule;
typedef refshim::ProgramGroup* OptixProgramGroup;
typedef refshim::Pipeline* OptixPipeline;
typedef refshim::Denoiser* OptixDenoiser;
typedef void* OptixGeometry;   // named only in dead declarations of the reference
typedef void* OptixMaterial;

typedef void (*OptixLogCallback)(unsigned int level, const char* tag, const char* message, void* cbdata);
struct OptixDeviceContextOptions { OptixLogCallback logCallbackFunction; void* logCallbackData; int logCallbackLevel; int validationMode; };

struct OptixAabb { float minX, minY, minZ, maxX, maxY, maxZ; };

enum OptixBuildFlags { OPTIX_BUILD_FLAG_NONE = 0, OPTIX_BUILD_FLAG_ALLOW_UPDATE = 1, OPTIX_BUILD_FLAG_ALLOW_COMPACTION = 2 };
enum OptixBuildOperation { OPTIX_BUILD_OPERATION_BUILD = 0x2161, OPTIX_BUILD_OPERATION_UPDATE = 0x2162 };
enum OptixMotionFlags { OPTIX_MOTION_FLAG_NONE = 0 };
enum OptixGeometryFlags { OPTIX_GEOMETRY_FLAG_NONE = 0, OPTIX_GEOMETRY_FLAG_DISABLE_ANYHIT = 1 };
enum OptixBuildInputType { OPTIX_BUILD_INPUT_TYPE_TRIANGLES = 0x2141, OPTIX_BUILD_INPUT_TYPE_CUSTOM_PRIMITIVES = 0x2142,
                           OPTIX_BUILD_INPUT_TYPE_INSTANCES = 0x2143 };
enum OptixInstanceFlags { OPTIX_INSTANCE_FLAG_NONE = 0 };
enum OptixAccelPropertyType { OPTIX_PROPERTY_TYPE_COMPACTED_SIZE = 0x2181 };
enum OptixTraversableType { OPTIX_TRAVERSABLE_TYPE_STATIC_TRANSFORM = 0x21C1, OPTIX_TRAVERSABLE_TYPE_MATRIX_MOTION_TRANSFORM = 0x21C2 };

struct OptixMotionOptions { unsigned short numKeys; unsigned short flags; float timeBegin; float timeEnd; };
struct OptixAccelBuildOptions { unsigned int buildFlags; OptixBuildOperation operation; OptixMotionOptions motionOptions; };
struct OptixBuildInputCustomPrimitiveArray {
    const CUdeviceptr* aabbBuffers; unsigned int numPrimitives; unsigned int strideInBytes; const unsigned int* flags;
    unsigned int numSbtRecords; CUdeviceptr sbtIndexOffsetBuffer; unsigned int sbtIndexOffsetSizeInBytes;
    unsigned int sbtIndexOffsetStrideInBytes; unsigned int primitiveIndexOffset;
};
struct OptixBuildInputInstanceArray { CUdeviceptr instances; unsigned int numInstances; unsigned int instanceStride; };
struct OptixBuildInput {
    OptixBuildInputType type;
    union { OptixBuildInputCustomPrimitiveArray customPrimitiveArray; OptixBuildInputInstanceArray instanceArray; };
};
struct OptixAccelBufferSizes { size_t outputSizeInBytes, tempSizeInBytes, tempUpdateSizeInBytes; };
struct OptixAccelEmitDesc { CUdeviceptr result; OptixAccelPropertyType type; };
struct OptixInstance {
    float transform[12]; unsigned int instanceId; unsigned int sbtOffset; unsigned int visibilityMask; unsigned int flags;
    OptixTraversableHandle traversableHandle; unsigned int pad[2];
};
struct OptixMatrixMotionTransform { OptixTraversableHandle child; OptixMotionOptions motionOptions; unsigned int pad[3]; float transform[2][12]; };

enum OptixCompileOptimizationLevel { OPTIX_COMPILE_OPTIMIZATION_DEFAULT = 0, OPTIX_COMPILE_OPTIMIZATION_LEVEL_0 = 0x2340 };
enum OptixCompileDebugLevel { OPTIX_COMPILE_DEBUG_LEVEL_DEFAULT = 0, OPTIX_COMPILE_DEBUG_LEVEL_FULL = 0x2352 };
enum OptixTraversableGraphFlags { OPTIX_TRAVERSABLE_GRAPH_FLAG_ALLOW_ANY = 0 };
enum OptixExceptionFlags { OPTIX_EXCEPTION_FLAG_NONE = 0, OPTIX_EXCEPTION_FLAG_STACK_OVERFLOW = 1 };
struct OptixModuleCompileOptions { int maxRegisterCount; OptixCompileOptimizationLevel optLevel; OptixCompileDebugLevel debugLevel; const void* boundValues; unsigned int numBoundValues; unsigned int numPayloadTypes; const void* payloadTypes; };
struct OptixPipelineCompileOptions {
    int usesMotionBlur; unsigned int traversableGraphFlags; int numPayloadValues; int numAttributeValues; unsigned int exceptionFlags;
    const char* pipelineLaunchParamsVariableName; unsigned int usesPrimitiveTypeFlags; int allowOpacityMicromaps;
};
struct OptixPipelineLinkOptions { unsigned int maxTraceDepth; };

enum OptixProgramGroupKind { OPTIX_PROGRAM_GROUP_KIND_RAYGEN = 0x2421, OPTIX_PROGRAM_GROUP_KIND_MISS = 0x2422, OPTIX_PROGRAM_GROUP_KIND_EXCEPTION = 0x2423,
                             OPTIX_PROGRAM_GROUP_KIND_HITGROUP = 0x2424, OPTIX_PROGRAM_GROUP_KIND_CALLABLES = 0x2425 };
enum OptixProgramGroupFlags { OPTIX_PROGRAM_GROUP_FLAGS_NONE = 0 };
struct OptixProgramGroupSingleModule { OptixModule module; const char* entryFunctionName; };
struct OptixProgramGroupHitgroup { OptixModule moduleCH; const char* entryFunctionNameCH; OptixModule moduleAH; const char* entryFunctionNameAH; OptixModule moduleIS; const char* entryFunctionNameIS; };
struct OptixProgramGroupCallables { OptixModule moduleDC; const char* entryFunctionNameDC; OptixModule moduleCC; const char* entryFunctionNameCC; };
struct OptixProgramGroupDesc {
    OptixProgramGroupKind kind; unsigned int flags;
    union { OptixProgramGroupSingleModule raygen; OptixProgramGroupSingleModule miss; OptixProgramGroupSingleModule exception;
            OptixProgramGroupCallables callables; OptixProgramGroupHitgroup hitgroup; };
};
struct OptixProgramGroupOptions { const void* payloadType; };
struct OptixStackSizes { unsigned int cssRG, cssMS, cssCH, cssAH, cssIS, cssCC, dssDC; };
struct OptixShaderBindingTable {
    CUdeviceptr raygenRecord; CUdeviceptr exceptionRecord;
    CUdeviceptr missRecordBase; unsigned int missRecordStrideInBytes; unsigned int missRecordCount;
    CUdeviceptr hitgroupRecordBase; unsigned int hitgroupRecordStrideInBytes; unsigned int hitgroupRecordCount;
    CUdeviceptr callablesRecordBase; unsigned int callablesRecordStrideInBytes; unsigned int callablesRecordCount;
};

enum OptixDenoiserModelKind { OPTIX_DENOISER_MODEL_KIND_LDR = 0x2322, OPTIX_DENOISER_MODEL_KIND_HDR = 0x2323 };
enum OptixPixelFormat { OPTIX_PIXEL_FORMAT_FLOAT3 = 0x2203, OPTIX_PIXEL_FORMAT_FLOAT4 = 0x2204 };
struct OptixDenoiserOptions { unsigned int guideAlbedo; unsigned int guideNormal; };
struct OptixDenoiserSizes { size_t stateSizeInBytes, withOverlapScratchSizeInBytes, withoutOverlapScratchSizeInBytes; unsigned int overlapWindowSizeInPixels;
                            size_t computeAverageColorSizeInBytes, computeIntensitySizeInBytes, internalGuideLayerPixelSizeInBytes; };
struct OptixDenoiserParams { int denoiseAlpha; CUdeviceptr hdrIntensity; float blendFactor; CUdeviceptr hdrAverageColor; unsigned int temporalModeUsePreviousLayers; };
struct OptixImage2D { CUdeviceptr data; unsigned int width, height, rowStrideInBytes, pixelStrideInBytes; OptixPixelFormat format; };
struct OptixDenoiserLayer { OptixImage2D input, previousOutput, output; int type; };
struct OptixDenoiserGuideLayer { OptixImage2D albedo, normal, flow, previousOutputInternalGuideLayer, outputInternalGuideLayer, flowTrustworthiness; };

// ---------------------------------------------------------------------------------------
// host API (implemented in ref_runtime.cpp)
// ---------------------------------------------------------------------------------------
OptixResult optixInit();
OptixResult optixDeviceContextCreate(CUcontext, const OptixDeviceContextOptions*, OptixDeviceContext* out);
OptixResult optixDeviceContextDestroy(OptixDeviceContext);
OptixResult optixModuleCreate(OptixDeviceContext, const OptixModuleCompileOptions*, const OptixPipelineCompileOptions*, const char* input,
                              size_t inputSize, char* log, size_t* logSize, OptixModule* out);
OptixResult optixModuleDestroy(OptixModule);
OptixResult optixProgramGroupCreate(OptixDeviceContext, const OptixProgramGroupDesc* descs, unsigned int n, const OptixProgramGroupOptions*,
                                    char* log, size_t* logSize, OptixProgramGroup* out);
OptixResult optixProgramGroupDestroy(OptixProgramGroup);
OptixResult optixPipelineCreate(OptixDeviceContext, const OptixPipelineCompileOptions*, const OptixPipelineLinkOptions*, const OptixProgramGroup*,
                                unsigned int n, char* log, size_t* logSize, OptixPipeline* out);
OptixResult optixPipelineDestroy(OptixPipeline);
OptixResult optixPipelineSetStackSize(OptixPipeline, unsigned int, unsigned int, unsigned int, unsigned int);
OptixResult optixUtilAccumulateStackSizes(OptixProgramGroup, OptixStackSizes*, OptixPipeline);
OptixResult optixUtilComputeStackSizes(const OptixStackSizes*, unsigned int, unsigned int, unsigned int, unsigned int*, unsigned int*, unsigned int*);
OptixResult optixSbtRecordPackHeader(OptixProgramGroup, void* sbtRecordHeaderHostPointer);
OptixResult optixAccelComputeMemoryUsage(OptixDeviceContext, const OptixAccelBuildOptions*, const OptixBuildInput*, unsigned int n, OptixAccelBufferSizes* out);
OptixResult optixAccelBuild(OptixDeviceContext, CUstream, const OptixAccelBuildOptions*, const OptixBuildInput*, unsigned int n, CUdeviceptr temp,
                            size_t tempSize, CUdeviceptr output, size_t outputSize, OptixTraversableHandle* out, const OptixAccelEmitDesc* emitted,
                            unsigned int nEmitted);
OptixResult optixAccelCompact(OptixDeviceContext, CUstream, OptixTraversableHandle in, CUdeviceptr output, size_t outputSize, OptixTraversableHandle* out);
OptixResult optixConvertPointerToTraversableHandle(OptixDeviceContext, CUdeviceptr pointer, OptixTraversableType, OptixTraversableHandle* out);
OptixResult optixLaunch(OptixPipeline, CUstream, CUdeviceptr params, size_t paramsSize, const OptixShaderBindingTable*, unsigned int w,
                        unsigned int h, unsigned int d);
OptixResult optixDenoiserCreate(OptixDeviceContext, OptixDenoiserModelKind, const OptixDenoiserOptions*, OptixDenoiser* out);
OptixResult optixDenoiserDestroy(OptixDenoiser);
OptixResult optixDenoiserComputeMemoryResources(OptixDenoiser, unsigned int w, unsigned int h, OptixDenoiserSizes* out);
OptixResult optixDenoiserSetup(OptixDenoiser, CUstream, unsigned int w, unsigned int h, CUdeviceptr state, size_t stateSize, CUdeviceptr scratch, size_t scratchSize);
OptixResult optixDenoiserComputeIntensity(OptixDenoiser, CUstream, const OptixImage2D* in, CUdeviceptr outIntensity, CUdeviceptr scratch, size_t scratchSize);
OptixResult optixDenoiserInvoke(OptixDenoiser, CUstream, const OptixDenoiserParams*, CUdeviceptr state, size_t stateSize, const OptixDenoiserGuideLayer*,
                                const OptixDenoiserLayer* layers, unsigned int nLayers, unsigned int offX, unsigned int offY, CUdeviceptr scratch,
                                size_t scratchSize);

// ---------------------------------------------------------------------------------------
// device API
// ---------------------------------------------------------------------------------------
enum OptixRayFlags { OPTIX_RAY_FLAG_NONE = 0, OPTIX_RAY_FLAG_DISABLE_ANYHIT = 1, OPTIX_RAY_FLAG_TERMINATE_ON_FIRST_HIT = 4,
                     OPTIX_RAY_FLAG_DISABLE_CLOSESTHIT = 8 };
enum OptixPayloadTypeID { OPTIX_PAYLOAD_TYPE_DEFAULT = 0, OPTIX_PAYLOAD_TYPE_ID_0 = 1 };

namespace refshim {
// What one host thread ("one GPU thread") sees while a program runs.
struct ThreadState {
    uint3 launch_index, launch_dim;
    // program context
    float3 world_o, world_d, obj_o, obj_d;
    float tmin, tmax, ray_time;
    unsigned int payload[2];
    unsigned int attr[8];
    const void* sbt_data;
    unsigned int instance_id;
    const float* o2w;   // 3x4 of the current instance, nullptr = identity
    const float* w2o;
    bool has_motion;
    float3 motion_t;    // interpolated translation of a matrix motion transform
    bool terminate_on_first_hit, done;
    const ProgramGroup* cur_pg;
    // outcome of the last optixTraverse (the "hit object")
    bool ho_hit;
    float ho_t;
    unsigned int ho_attr[8];
    const void* ho_sbt_data;
    unsigned int ho_instance_id;
    const ProgramGroup* ho_pg;
    float3 ho_world_o, ho_world_d;
    float ho_tmin;
    unsigned int ho_miss_index;
};
extern thread_local ThreadState ts;
void* callable_entry(unsigned int sbtIndex);   // direct callable at callablesRecordBase + sbtIndex * stride
void traverse(OptixTraversableHandle handle, float3 o, float3 d, float tmin, float tmax, float rayTime, unsigned int rayFlags,
              unsigned int sbtOffset, unsigned int sbtStride, unsigned int missIndex, unsigned int p0, unsigned int p1);
void invoke(unsigned int p0, unsigned int p1);
bool report_intersection(float t, unsigned int kind, const unsigned int* a);
float4 tex2d(cudaTextureObject_t tex, float u, float v);
}  // namespace refshim

static inline uint3 optixGetLaunchIndex() { return refshim::ts.launch_index; }
static inline uint3 optixGetLaunchDimensions() { return refshim::ts.launch_dim; }
static inline unsigned int optixGetPayload_0() { return refshim::ts.payload[0]; }
static inline unsigned int optixGetPayload_1() { return refshim::ts.payload[1]; }
static inline unsigned int optixGetAttribute_0() { return refshim::ts.attr[0]; }
static inline unsigned int optixGetAttribute_1() { return refshim::ts.attr[1]; }
static inline unsigned int optixGetAttribute_2() { return refshim::ts.attr[2]; }
static inline unsigned int optixGetAttribute_3() { return refshim::ts.attr[3]; }
static inline unsigned int optixGetAttribute_4() { return refshim::ts.attr[4]; }
static inline unsigned int optixGetAttribute_5() { return refshim::ts.attr[5]; }
static inline unsigned int optixGetAttribute_6() { return refshim::ts.attr[6]; }
static inline unsigned int optixGetAttribute_7() { return refshim::ts.attr[7]; }
static inline unsigned int optixGetInstanceId() { return refshim::ts.instance_id; }
static inline CUdeviceptr optixGetSbtDataPointer() { return (CUdeviceptr)refshim::ts.sbt_data; }
static inline float3 optixGetObjectRayOrigin() { return refshim::ts.obj_o; }
static inline float3 optixGetObjectRayDirection() { return refshim::ts.obj_d; }
static inline float3 optixGetWorldRayOrigin() { return refshim::ts.world_o; }
static inline float3 optixGetWorldRayDirection() { return refshim::ts.world_d; }
static inline float optixGetRayTmin() { return refshim::ts.tmin; }
static inline float optixGetRayTmax() { return refshim::ts.tmax; }
static inline float optixGetRayTime() { return refshim::ts.ray_time; }

// point: M * (p, 1) with the row evaluated left to right; normal: W2O^T * n (inverse transpose)
static inline float3 optixTransformPointFromObjectToWorldSpace(float3 p) {
    const refshim::ThreadState& s = refshim::ts;
    if (s.has_motion) return make_float3(p.x + s.motion_t.x, p.y + s.motion_t.y, p.z + s.motion_t.z);
    const float* m = s.o2w;
    if (!m) return p;
    return make_float3(m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3], m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7],
                       m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]);
}
static inline float3 optixTransformNormalFromObjectToWorldSpace(float3 n) {
    const refshim::ThreadState& s = refshim::ts;
    const float* m = s.w2o;
    if (s.has_motion || !m) return n;
    return make_float3(m[0] * n.x + m[4] * n.y + m[8] * n.z, m[1] * n.x + m[5] * n.y + m[9] * n.z, m[2] * n.x + m[6] * n.y + m[10] * n.z);
}

static inline bool optixReportIntersection(float t, unsigned int kind, unsigned int a0, unsigned int a1, unsigned int a2, unsigned int a3,
                                           unsigned int a4, unsigned int a5, unsigned int a6, unsigned int a7) {
    const unsigned int a[8] = {a0, a1, a2, a3, a4, a5, a6, a7};
    return refshim::report_intersection(t, kind, a);
}
static inline void optixTraverse(OptixPayloadTypeID, OptixTraversableHandle handle, float3 o, float3 d, float tmin, float tmax, float rayTime,
                                 OptixVisibilityMask, unsigned int rayFlags, unsigned int sbtOffset, unsigned int sbtStride, unsigned int missIndex,
                                 unsigned int& p0, unsigned int& p1) {
    refshim::traverse(handle, o, d, tmin, tmax, rayTime, rayFlags, sbtOffset, sbtStride, missIndex, p0, p1);
}
static inline void optixTraverse(OptixTraversableHandle handle, float3 o, float3 d, float tmin, float tmax, float rayTime, OptixVisibilityMask,
                                 unsigned int rayFlags, unsigned int sbtOffset, unsigned int sbtStride, unsigned int missIndex, unsigned int& p0,
                                 unsigned int& p1) {
    refshim::traverse(handle, o, d, tmin, tmax, rayTime, rayFlags, sbtOffset, sbtStride, missIndex, p0, p1);
}
static inline void optixReorder() {}
static inline void optixInvoke(OptixPayloadTypeID, unsigned int& p0, unsigned int& p1) { refshim::invoke(p0, p1); }
static inline bool optixHitObjectIsHit() { return refshim::ts.ho_hit; }

template <typename ReturnT, typename... ArgTypes>
static inline ReturnT optixDirectCall(unsigned int sbtIndex, ArgTypes... args) {
    typedef ReturnT (*Fn)(ArgTypes...);
    return ((Fn)refshim::callable_entry(sbtIndex))(args...);
}

// texture fetch (texture_indirect_functions.h in the CUDA toolkit)
template <typename T>
static inline T tex2D(cudaTextureObject_t tex, float u, float v);
template <>
inline float4 tex2D<float4>(cudaTextureObject_t tex, float u, float v) { return refshim::tex2d(tex, u, v); }
