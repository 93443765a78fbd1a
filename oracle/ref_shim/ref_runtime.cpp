// ORACLE — TEST INFRASTRUCTURE ONLY (oracle/ref_shim).
//
// ref_runtime.cpp — the implementation behind oracle/ref_shim/optix.h and cuda_runtime.h:
// "device memory" is host memory, an acceleration structure is a list, optixLaunch is a
// loop over pixels on host threads, traversal is a loop over the instances calling the
// reference's own __intersection__ programs, and the shader binding table is followed the
// way OptiX documents it (record = base + (instance.sbtOffset + rayTypeOffset) * stride,
// header = the program group).  The rules stated at the top of optix.h are implemented here.
#include "optix.h"
#include "cuda/random.h"

#include <atomic>
#include <cfloat>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

// ---- the reference's programs (defined in its .cu files; only their addresses are taken) ----
#define REF_ENTRY_LIST(X)                                                                                          \
    X(__raygen__Program) X(__miss__Program) X(__closesthit__radiance) X(__intersection__sphere)                   \
    X(__intersection__movingsphere) X(__intersection__hitRectX) X(__intersection__hitRectY)                       \
    X(__intersection__hitRectZ) X(__intersection__hitVolumebox) X(__intersection__hitVolumesphere)                \
    X(__direct_callable__perspectiveCamera) X(__direct_callable__diffuselight) X(__direct_callable__dielectricMat) \
    X(__direct_callable__ISOTROPIC) X(__direct_callable__lambertian) X(__direct_callable__metal)                  \
    X(__direct_callable__normal) X(__direct_callable__checkerTexture) X(__direct_callable__constantTexture)       \
    X(__direct_callable__imagetexture) X(__direct_callable__noisetexture) X(__direct_callable__nullTexture)       \
    X(__direct_callable__cosineGenerate) X(__direct_callable__mixtureBIAS_generate)                               \
    X(__direct_callable__mixture_generate) X(__direct_callable__rect_x_generate)                                  \
    X(__direct_callable__rect_y_generate) X(__direct_callable__rect_z_generate) X(__direct_callable__cosineValue) \
    X(__direct_callable__mixtureBIAS_value) X(__direct_callable__mixture_value) X(__direct_callable__rect_x_value) \
    X(__direct_callable__rect_y_value) X(__direct_callable__rect_z_value)                                         \
    X(__direct_callable__eval_bsdf_diffuse_reflection) X(__direct_callable__eval_bsdf_dielect_reflection)         \
    X(__direct_callable__eval_bsdf_metal_reflection)
extern "C" {
#define X(name) void name(void);
REF_ENTRY_LIST(X)
#undef X
}

namespace refshim {

// provided by ref_glue.cpp (the one stand-in file that sees the reference's own types)
void* params_storage();
unsigned int* prd_seed(unsigned int p0, unsigned int p1);

typedef void (*Fn)(void);
static Fn lookup_entry(const char* name) {
    if (!name) return nullptr;
#define X(n) if (std::strcmp(name, #n) == 0) return (Fn)&n;
    REF_ENTRY_LIST(X)
#undef X
    std::fprintf(stderr, "ref_shim: unknown program entry '%s'\n", name);
    std::abort();
}

struct Context { int unused; };
struct Module { int unused; };
struct Pipeline { int unused; };
struct Denoiser { int unused; };
struct ProgramGroup {
    OptixProgramGroupKind kind;
    Fn rg = nullptr, ms = nullptr, ch = nullptr, is = nullptr, dc = nullptr;
    std::string name;   // entry name: raygen / miss / intersection / direct callable
    bool is_volume = false;
};

enum AccelKind { ACCEL_GAS = 0x6a51, ACCEL_IAS = 0x6a52, ACCEL_MOTION = 0x6a53 };
struct Accel { int kind; };
struct Gas : Accel { std::vector<OptixAabb> aabbs; };
struct InstanceRec { OptixInstance inst; bool identity; float w2o[12]; };
struct Ias : Accel { std::vector<InstanceRec> instances; };
struct Motion : Accel { const OptixMatrixMotionTransform* mt; };
struct TexObj { cudaArray_t array; };

thread_local ThreadState ts;
static thread_local unsigned long long tl_traversals = 0;

// launch-wide state
static OptixShaderBindingTable g_sbt;
static int g_threads = 0;
static int g_q23_normalise = 0;
static std::atomic<unsigned long long> g_traversals{0}, g_launches{0};
static OptixTraversableHandle g_last_ias = 0;

int set_threads(int n) { int old = g_threads; g_threads = n; return old; }
int set_q23_normalise(int on) { int old = g_q23_normalise; g_q23_normalise = on; return old; }
unsigned long long traversal_count() { return g_traversals.load(); }
const OptixShaderBindingTable& last_sbt() { return g_sbt; }
OptixTraversableHandle last_ias() { return g_last_ias; }

// ---- inspection helpers for ref_api.cpp ----
int ias_size(OptixTraversableHandle h) {
    const Accel* a = (const Accel*)h;
    return (a && a->kind == ACCEL_IAS) ? (int)((const Ias*)a)->instances.size() : -1;
}
const OptixInstance* ias_instance(OptixTraversableHandle h, int i, const float** w2o, int* identity) {
    const Ias* a = (const Ias*)h;
    if (w2o) *w2o = a->instances[i].w2o;
    if (identity) *identity = a->instances[i].identity;
    return &a->instances[i].inst;
}
// child of an instance: 0 = GAS, 1 = matrix motion transform over a GAS
int child_info(OptixTraversableHandle child, OptixAabb* aabb, const OptixMatrixMotionTransform** mt) {
    const Accel* a = (const Accel*)child;
    int kind = 0;
    if (a->kind == ACCEL_MOTION) {
        kind = 1;
        if (mt) *mt = ((const Motion*)a)->mt;
        a = (const Accel*)((const Motion*)a)->mt->child;
    }
    if (aabb && a && a->kind == ACCEL_GAS && !((const Gas*)a)->aabbs.empty()) *aabb = ((const Gas*)a)->aabbs[0];
    return kind;
}
const char* program_name(const void* sbtRecordHeader) {
    const ProgramGroup* pg;
    std::memcpy(&pg, sbtRecordHeader, sizeof(pg));
    return pg ? pg->name.c_str() : "";
}
const cudaArray* texture_array(cudaTextureObject_t t) { return t ? ((const TexObj*)t)->array : nullptr; }

// ---- SBT access ----
static inline const ProgramGroup* record_pg(CUdeviceptr base, unsigned int stride, unsigned int index) {
    const ProgramGroup* pg;
    std::memcpy(&pg, (const char*)base + (size_t)index * stride, sizeof(pg));
    return pg;
}
void* callable_entry(unsigned int sbtIndex) {
    if (sbtIndex >= g_sbt.callablesRecordCount) {
        std::fprintf(stderr, "ref_shim: optixDirectCall index %u outside the %u callable records\n", sbtIndex, g_sbt.callablesRecordCount);
        std::abort();
    }
    return (void*)record_pg(g_sbt.callablesRecordBase, g_sbt.callablesRecordStrideInBytes, sbtIndex)->dc;
}

// ---- traversal ----
bool report_intersection(float t, unsigned int, const unsigned int* a) {
    ThreadState& s = ts;
    if (!(t >= s.tmin && t <= s.tmax)) return false;
    s.tmax = t;
    s.ho_hit = true;
    s.ho_t = t;
    for (int i = 0; i < 8; i++) s.ho_attr[i] = a[i];
    s.ho_sbt_data = s.sbt_data;
    s.ho_instance_id = s.instance_id;
    s.ho_pg = s.cur_pg;
    if (s.terminate_on_first_hit) s.done = true;
    return true;
}

void traverse(OptixTraversableHandle handle, float3 o, float3 d, float tmin, float tmax, float rayTime, unsigned int rayFlags,
              unsigned int sbtOffset, unsigned int sbtStride, unsigned int missIndex, unsigned int p0, unsigned int p1) {
    tl_traversals++;
    const ThreadState saved = ts;
    ThreadState& s = ts;
    s.world_o = o; s.world_d = d; s.tmin = tmin; s.tmax = tmax; s.ray_time = rayTime;
    s.payload[0] = p0; s.payload[1] = p1;
    s.terminate_on_first_hit = (rayFlags & OPTIX_RAY_FLAG_TERMINATE_ON_FIRST_HIT) != 0;
    s.done = false;
    s.ho_hit = false;
    const Ias* ias = (const Ias*)handle;
    if (!ias || ias->kind != ACCEL_IAS) { std::fprintf(stderr, "ref_shim: optixTraverse on a handle that is not an IAS\n"); std::abort(); }
    unsigned int* seedp = g_q23_normalise ? prd_seed(p0, p1) : nullptr;
    const unsigned int seed0 = seedp ? *seedp : 0u;
    for (int i = (int)ias->instances.size() - 1; i >= 0 && !s.done; i--) {
        const InstanceRec& ir = ias->instances[i];
        const unsigned int rec = ir.inst.sbtOffset + sbtOffset;   // one build input, one record per GAS: GAS index 0
        (void)sbtStride;
        if (rec >= g_sbt.hitgroupRecordCount) {
            std::fprintf(stderr, "ref_shim: instance %d: SBT index %u outside the %u hit records\n", i, rec, g_sbt.hitgroupRecordCount);
            std::abort();
        }
        const char* record = (const char*)g_sbt.hitgroupRecordBase + (size_t)rec * g_sbt.hitgroupRecordStrideInBytes;
        const ProgramGroup* pg;
        std::memcpy(&pg, record, sizeof(pg));
        if (!pg || !pg->is) continue;
        s.sbt_data = record + OPTIX_SBT_RECORD_HEADER_SIZE;
        s.instance_id = ir.inst.instanceId;
        // world -> object
        float3 oo = o, od = d;
        s.o2w = nullptr; s.w2o = nullptr; s.has_motion = false;
        if (!ir.identity) {
            const float* m = ir.w2o;
            oo = make_float3(m[0] * o.x + m[1] * o.y + m[2] * o.z + m[3], m[4] * o.x + m[5] * o.y + m[6] * o.z + m[7],
                             m[8] * o.x + m[9] * o.y + m[10] * o.z + m[11]);
            od = make_float3(m[0] * d.x + m[1] * d.y + m[2] * d.z, m[4] * d.x + m[5] * d.y + m[6] * d.z, m[8] * d.x + m[9] * d.y + m[10] * d.z);
            s.o2w = ir.inst.transform;
            s.w2o = ir.w2o;
        }
        const Accel* child = (const Accel*)ir.inst.traversableHandle;
        if (child && child->kind == ACCEL_MOTION) {
            const OptixMatrixMotionTransform* mt = ((const Motion*)child)->mt;
            float f = (rayTime - mt->motionOptions.timeBegin) / (mt->motionOptions.timeEnd - mt->motionOptions.timeBegin);
            if (f < 0.f) f = 0.f;
            if (f > 1.f) f = 1.f;
            const float* k0 = mt->transform[0];
            const float* k1 = mt->transform[1];
            s.motion_t = make_float3(k0[3] * (1.0f - f) + k1[3] * f, k0[7] * (1.0f - f) + k1[7] * f, k0[11] * (1.0f - f) + k1[11] * f);
            s.has_motion = true;
            oo = make_float3(oo.x - s.motion_t.x, oo.y - s.motion_t.y, oo.z - s.motion_t.z);
        }
        s.obj_o = oo; s.obj_d = od;
        s.cur_pg = pg;
        // Q23: the constant-medium programs draw their free-flight distance from the
        // path seed INSIDE the intersection program, so the reference's image depends on the driver's
        // traversal order.  With q23_normalise the draw comes from tea<4>(path seed at ray start,
        // instance index) and the path seed is left alone — the rule oracle/oracle.cpp states.
        if (seedp && pg->is_volume) *seedp = tea<4>(seed0, (unsigned int)i);
        pg->is();
        if (seedp && pg->is_volume) *seedp = seed0;
    }
    // the program context of the caller comes back; only the hit object is new
    ThreadState out = saved;
    out.ho_hit = s.ho_hit;
    out.ho_t = s.ho_t;
    for (int i = 0; i < 8; i++) out.ho_attr[i] = s.ho_attr[i];
    out.ho_sbt_data = s.ho_sbt_data;
    out.ho_instance_id = s.ho_instance_id;
    out.ho_pg = s.ho_pg;
    out.ho_world_o = o; out.ho_world_d = d;
    out.ho_tmin = tmin;
    out.ho_miss_index = missIndex;
    ts = out;
}

void invoke(unsigned int p0, unsigned int p1) {
    const ThreadState saved = ts;
    ThreadState& s = ts;
    s.payload[0] = p0; s.payload[1] = p1;
    s.world_o = saved.ho_world_o; s.world_d = saved.ho_world_d; s.tmin = saved.ho_tmin;
    if (saved.ho_hit) {
        s.tmax = saved.ho_t;
        for (int i = 0; i < 8; i++) s.attr[i] = saved.ho_attr[i];
        s.sbt_data = saved.ho_sbt_data;
        s.instance_id = saved.ho_instance_id;
        saved.ho_pg->ch();
    } else {
        record_pg(g_sbt.missRecordBase, g_sbt.missRecordStrideInBytes, saved.ho_miss_index)->ms();
    }
    ts = saved;
}

// ---- texture unit: normalised coordinates, clamp, fp32 bilinear on RGBA8 ----
float4 tex2d(cudaTextureObject_t tex, float u, float v) {
    const cudaArray* a = ((const TexObj*)tex)->array;
    const int W = a->width, H = a->height;
    float x = u * (float)W - 0.5f, y = v * (float)H - 0.5f;
    if (x != x) x = 0.f;
    if (y != y) y = 0.f;
    const float xf = floorf(x), yf = floorf(y);
    const float ax = x - xf, ay = y - yf;
    auto cl = [](float c, int n) { float hi = (float)(n - 1); c = c < 0.f ? 0.f : c; c = c > hi ? hi : c; return (int)c; };
    const int x0 = cl(xf, W), x1 = cl(xf + 1.f, W), y0 = cl(yf, H), y1 = cl(yf + 1.f, H);
    auto px = [&](int xi, int yi, int c) { return (float)a->rgba[((size_t)yi * W + xi) * 4 + c] / 255.0f; };
    float out[4];
    for (int c = 0; c < 4; c++) {
        const float top = px(x0, y0, c) * (1.0f - ax) + px(x1, y0, c) * ax;
        const float bot = px(x0, y1, c) * (1.0f - ax) + px(x1, y1, c) * ax;
        out[c] = top * (1.0f - ay) + bot * ay;
    }
    return make_float4(out[0], out[1], out[2], out[3]);
}

}  // namespace refshim
using namespace refshim;

// =========================================================================================
// CUDA runtime stand-ins
// =========================================================================================
cudaError_t cudaMalloc(void** p, size_t n) { *p = std::calloc(n ? n : 1, 1); return *p ? cudaSuccess : cudaErrorInvalidValue; }
cudaError_t cudaFree(void*) { return cudaSuccess; }   // the reference frees some buffers twice / never: keep everything alive
cudaError_t cudaMemcpy(void* dst, const void* src, size_t n, cudaMemcpyKind) { if (n) std::memcpy(dst, src, n); return cudaSuccess; }
cudaError_t cudaStreamCreate(cudaStream_t* s) { *s = nullptr; return cudaSuccess; }
cudaError_t cudaStreamDestroy(cudaStream_t) { return cudaSuccess; }
cudaError_t cudaSetDevice(int) { return cudaSuccess; }
cudaError_t cudaMallocArray(cudaArray_t* a, const cudaChannelFormatDesc*, size_t w, size_t h) {
    *a = new cudaArray{(int)w, (int)h, (unsigned char*)std::calloc(w * h, 4)};
    return cudaSuccess;
}
cudaError_t cudaMemcpy2DToArray(cudaArray_t dst, size_t, size_t, const void* src, size_t spitch, size_t widthBytes, size_t height, cudaMemcpyKind) {
    for (size_t r = 0; r < height; r++) std::memcpy(dst->rgba + r * (size_t)dst->width * 4, (const char*)src + r * spitch, widthBytes);
    return cudaSuccess;
}
cudaError_t cudaCreateTextureObject(cudaTextureObject_t* out, const cudaResourceDesc* res, const cudaTextureDesc*, const cudaResourceViewDesc*) {
    *out = (cudaTextureObject_t) new TexObj{res->res.array.array};
    return cudaSuccess;
}

// =========================================================================================
// OptiX host API stand-ins
// =========================================================================================
OptixResult optixInit() { return OPTIX_SUCCESS; }
OptixResult optixDeviceContextCreate(CUcontext, const OptixDeviceContextOptions*, OptixDeviceContext* out) { *out = new Context{0}; return OPTIX_SUCCESS; }
OptixResult optixDeviceContextDestroy(OptixDeviceContext) { return OPTIX_SUCCESS; }
OptixResult optixModuleCreate(OptixDeviceContext, const OptixModuleCompileOptions*, const OptixPipelineCompileOptions*, const char*, size_t, char*,
                              size_t*, OptixModule* out) { *out = new Module{0}; return OPTIX_SUCCESS; }
OptixResult optixModuleDestroy(OptixModule) { return OPTIX_SUCCESS; }
OptixResult optixProgramGroupCreate(OptixDeviceContext, const OptixProgramGroupDesc* descs, unsigned int n, const OptixProgramGroupOptions*, char*,
                                    size_t*, OptixProgramGroup* out) {
    for (unsigned int i = 0; i < n; i++) {
        ProgramGroup* pg = new ProgramGroup();
        pg->kind = descs[i].kind;
        switch (descs[i].kind) {
        case OPTIX_PROGRAM_GROUP_KIND_RAYGEN: pg->rg = lookup_entry(descs[i].raygen.entryFunctionName); pg->name = descs[i].raygen.entryFunctionName; break;
        case OPTIX_PROGRAM_GROUP_KIND_MISS: pg->ms = lookup_entry(descs[i].miss.entryFunctionName); pg->name = descs[i].miss.entryFunctionName; break;
        case OPTIX_PROGRAM_GROUP_KIND_HITGROUP:
            pg->ch = lookup_entry(descs[i].hitgroup.entryFunctionNameCH);
            pg->is = lookup_entry(descs[i].hitgroup.entryFunctionNameIS);
            pg->name = descs[i].hitgroup.entryFunctionNameIS ? descs[i].hitgroup.entryFunctionNameIS : "";
            pg->is_volume = pg->name.find("Volume") != std::string::npos;
            break;
        case OPTIX_PROGRAM_GROUP_KIND_CALLABLES: pg->dc = lookup_entry(descs[i].callables.entryFunctionNameDC); pg->name = descs[i].callables.entryFunctionNameDC; break;
        default: break;
        }
        out[i] = pg;
    }
    return OPTIX_SUCCESS;
}
OptixResult optixProgramGroupDestroy(OptixProgramGroup) { return OPTIX_SUCCESS; }
OptixResult optixPipelineCreate(OptixDeviceContext, const OptixPipelineCompileOptions*, const OptixPipelineLinkOptions*, const OptixProgramGroup*,
                                unsigned int, char*, size_t*, OptixPipeline* out) { *out = new Pipeline{0}; return OPTIX_SUCCESS; }
OptixResult optixPipelineDestroy(OptixPipeline) { return OPTIX_SUCCESS; }
OptixResult optixPipelineSetStackSize(OptixPipeline, unsigned int, unsigned int, unsigned int, unsigned int) { return OPTIX_SUCCESS; }
OptixResult optixUtilAccumulateStackSizes(OptixProgramGroup, OptixStackSizes*, OptixPipeline) { return OPTIX_SUCCESS; }
OptixResult optixUtilComputeStackSizes(const OptixStackSizes*, unsigned int, unsigned int, unsigned int, unsigned int* a, unsigned int* b, unsigned int* c) {
    *a = *b = *c = 0;
    return OPTIX_SUCCESS;
}
OptixResult optixSbtRecordPackHeader(OptixProgramGroup pg, void* header) {
    std::memset(header, 0, OPTIX_SBT_RECORD_HEADER_SIZE);
    std::memcpy(header, &pg, sizeof(pg));
    return OPTIX_SUCCESS;
}
OptixResult optixAccelComputeMemoryUsage(OptixDeviceContext, const OptixAccelBuildOptions*, const OptixBuildInput*, unsigned int, OptixAccelBufferSizes* out) {
    out->outputSizeInBytes = 256; out->tempSizeInBytes = 256; out->tempUpdateSizeInBytes = 0;
    return OPTIX_SUCCESS;
}
static void invert_3x4(const float* m, float* out) {   // double cofactors, rounded once
    double a[3][3], t[3];
    for (int i = 0; i < 3; i++) { for (int j = 0; j < 3; j++) a[i][j] = m[i * 4 + j]; t[i] = m[i * 4 + 3]; }
    const double c00 = a[1][1] * a[2][2] - a[1][2] * a[2][1];
    const double c01 = a[1][2] * a[2][0] - a[1][0] * a[2][2];
    const double c02 = a[1][0] * a[2][1] - a[1][1] * a[2][0];
    const double det = a[0][0] * c00 + a[0][1] * c01 + a[0][2] * c02;
    double inv[3][3];
    inv[0][0] = c00 / det; inv[1][0] = c01 / det; inv[2][0] = c02 / det;
    inv[0][1] = (a[0][2] * a[2][1] - a[0][1] * a[2][2]) / det;
    inv[1][1] = (a[0][0] * a[2][2] - a[0][2] * a[2][0]) / det;
    inv[2][1] = (a[0][1] * a[2][0] - a[0][0] * a[2][1]) / det;
    inv[0][2] = (a[0][1] * a[1][2] - a[0][2] * a[1][1]) / det;
    inv[1][2] = (a[0][2] * a[1][0] - a[0][0] * a[1][2]) / det;
    inv[2][2] = (a[0][0] * a[1][1] - a[0][1] * a[1][0]) / det;
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) out[i * 4 + j] = (float)inv[i][j];
        out[i * 4 + 3] = (float)(-(inv[i][0] * t[0] + inv[i][1] * t[1] + inv[i][2] * t[2]));
    }
}
OptixResult optixAccelBuild(OptixDeviceContext, CUstream, const OptixAccelBuildOptions*, const OptixBuildInput* in, unsigned int, CUdeviceptr, size_t,
                            CUdeviceptr, size_t outputSize, OptixTraversableHandle* out, const OptixAccelEmitDesc* emitted, unsigned int nEmitted) {
    if (in->type == OPTIX_BUILD_INPUT_TYPE_CUSTOM_PRIMITIVES) {
        Gas* g = new Gas();
        g->kind = ACCEL_GAS;
        const OptixAabb* boxes = (const OptixAabb*)in->customPrimitiveArray.aabbBuffers[0];
        g->aabbs.assign(boxes, boxes + in->customPrimitiveArray.numPrimitives);
        *out = (OptixTraversableHandle)(Accel*)g;
    } else if (in->type == OPTIX_BUILD_INPUT_TYPE_INSTANCES) {
        Ias* a = new Ias();
        a->kind = ACCEL_IAS;
        const OptixInstance* inst = (const OptixInstance*)in->instanceArray.instances;
        static const float ident[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
        for (unsigned int i = 0; i < in->instanceArray.numInstances; i++) {
            InstanceRec r;
            r.inst = inst[i];
            r.identity = std::memcmp(inst[i].transform, ident, sizeof(ident)) == 0;
            invert_3x4(inst[i].transform, r.w2o);
            a->instances.push_back(r);
        }
        *out = (OptixTraversableHandle)(Accel*)a;
        g_last_ias = *out;
    } else {
        return OPTIX_ERROR_INVALID_VALUE;
    }
    for (unsigned int i = 0; i < nEmitted; i++) *(size_t*)emitted[i].result = outputSize;   // "compaction saves nothing"
    return OPTIX_SUCCESS;
}
OptixResult optixAccelCompact(OptixDeviceContext, CUstream, OptixTraversableHandle in, CUdeviceptr, size_t, OptixTraversableHandle* out) { *out = in; return OPTIX_SUCCESS; }
OptixResult optixConvertPointerToTraversableHandle(OptixDeviceContext, CUdeviceptr pointer, OptixTraversableType type, OptixTraversableHandle* out) {
    if (type != OPTIX_TRAVERSABLE_TYPE_MATRIX_MOTION_TRANSFORM) return OPTIX_ERROR_INVALID_VALUE;
    Motion* m = new Motion();
    m->kind = ACCEL_MOTION;
    m->mt = (const OptixMatrixMotionTransform*)pointer;
    static const float rot[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    for (int k = 0; k < 2; k++)
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++)
                if (m->mt->transform[k][i * 4 + j] != rot[i * 3 + j]) {
                    std::fprintf(stderr, "ref_shim: motion keys with a rotation/scale part are not supported\n");
                    return OPTIX_ERROR_INVALID_VALUE;
                }
    *out = (OptixTraversableHandle)(Accel*)m;
    return OPTIX_SUCCESS;
}

OptixResult optixLaunch(OptixPipeline, CUstream, CUdeviceptr params, size_t paramsSize, const OptixShaderBindingTable* sbt, unsigned int w,
                        unsigned int h, unsigned int) {
    std::memcpy(params_storage(), (const void*)params, paramsSize);
    g_sbt = *sbt;
    g_launches.fetch_add(1);
    const ProgramGroup* rg = record_pg(sbt->raygenRecord, 0, 0);
    if (std::getenv("REF_SKIP_LAUNCH")) return OPTIX_SUCCESS;   // CLI tests: parse, build the scene, print a black frame
    int nt = g_threads;
    if (nt <= 0) { const char* e = std::getenv("REF_THREADS"); nt = e ? std::atoi(e) : (int)std::thread::hardware_concurrency(); }
    if (nt <= 0) nt = 1;
    if (!g_q23_normalise) { const char* e = std::getenv("REF_Q23_NORMALISE"); if (e && std::atoi(e)) g_q23_normalise = 1; }
    std::atomic<unsigned int> next_row(0);
    auto worker = [&]() {
        ts = ThreadState{};
        ts.launch_dim = make_uint3(w, h, 1);
        for (;;) {
            const unsigned int y = next_row.fetch_add(1);
            if (y >= h) break;
            for (unsigned int x = 0; x < w; x++) {
                ts.launch_index = make_uint3(x, y, 0);
                rg->rg();
            }
        }
        g_traversals.fetch_add(tl_traversals);
        tl_traversals = 0;
    };
    if (nt == 1) { worker(); return OPTIX_SUCCESS; }
    std::vector<std::thread> th;
    for (int i = 0; i < nt; i++) th.emplace_back(worker);
    for (auto& t : th) t.join();
    return OPTIX_SUCCESS;
}

OptixResult optixDenoiserCreate(OptixDeviceContext, OptixDenoiserModelKind, const OptixDenoiserOptions*, OptixDenoiser* out) { *out = new Denoiser{0}; return OPTIX_SUCCESS; }
OptixResult optixDenoiserDestroy(OptixDenoiser) { return OPTIX_SUCCESS; }
OptixResult optixDenoiserComputeMemoryResources(OptixDenoiser, unsigned int, unsigned int, OptixDenoiserSizes* out) {
    std::memset(out, 0, sizeof(*out));
    out->stateSizeInBytes = 16; out->withOverlapScratchSizeInBytes = 16; out->withoutOverlapScratchSizeInBytes = 16;
    return OPTIX_SUCCESS;
}
OptixResult optixDenoiserSetup(OptixDenoiser, CUstream, unsigned int, unsigned int, CUdeviceptr, size_t, CUdeviceptr, size_t) { return OPTIX_SUCCESS; }
OptixResult optixDenoiserComputeIntensity(OptixDenoiser, CUstream, const OptixImage2D*, CUdeviceptr, CUdeviceptr, size_t) { return OPTIX_SUCCESS; }
// identity: the pre-denoise frame is the parity target
OptixResult optixDenoiserInvoke(OptixDenoiser, CUstream, const OptixDenoiserParams*, CUdeviceptr, size_t, const OptixDenoiserGuideLayer*,
                                const OptixDenoiserLayer* layers, unsigned int, unsigned int, unsigned int, CUdeviceptr, size_t) {
    const OptixImage2D& in = layers[0].input;
    std::memcpy((void*)layers[0].output.data, (const void*)in.data, (size_t)in.height * in.rowStrideInBytes);
    return OPTIX_SUCCESS;
}
