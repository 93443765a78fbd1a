// ORACLE — TEST INFRASTRUCTURE ONLY (oracle/ref_shim).
//
// ref_api.cpp — C entry points of oracle/_ref/libref.so (libm transcendentals) and
// oracle/_ref/libref_dm.so (the same sources with sinf/cosf/logf/powf/atan2f/asinf of the
// DEVICE files mapped to oracle/detmath.h).  Both contain the reference's own main.cpp,
// Director.cpp, scene/geometry/material/texture headers and all 26 device .cu files, compiled
// where they lie under /root/reference against the stand-in headers of this directory.
//
//   ref_render_scene   Director::init -> createScene(n) -> renderFrame(), exactly the calls of
//                      RestOfLife/main.cpp:148-157, and a copy of the frame buffer the launch wrote
//   ref_scene_desc     the state that launch ran on (IAS instances, SBT records, launch
//                      parameters) flattened to the C-ABI's rtw_scene_desc, so that the product's
//                      host scene builder can be compared with what the reference really built
//   ref_render_desc    the reference's DEVICE programs launched on an arbitrary rtw_scene_desc
//                      (IAS / SBT / launch parameters assembled here the way Director.cpp:483-553,
//                      628-885 assembles them), for branches no built-in scene reaches
//   ref_main           the reference's main() (compiled with -Dmain=ref_main)
#include <unistd.h>
#include <fcntl.h>

#include <map>
#include <string>
#include <vector>

#include "Director.h"
#include "../../include/rtw.h"

namespace refshim {
int set_threads(int n);
int set_q23_normalise(int on);
unsigned long long traversal_count();
const OptixShaderBindingTable& last_sbt();
OptixTraversableHandle last_ias();
int ias_size(OptixTraversableHandle h);
const OptixInstance* ias_instance(OptixTraversableHandle h, int i, const float** w2o, int* identity);
int child_info(OptixTraversableHandle child, OptixAabb* aabb, const OptixMatrixMotionTransform** mt);
const char* program_name(const void* sbtRecordHeader);
const cudaArray* texture_array(cudaTextureObject_t t);
void* params_storage();
}
extern "C" SysParamter Parameter;

namespace {

struct QuietStderr {   // Director::createModule reports 26 missing .optixir files (Director.cpp:154-181)
    int saved = -1;
    explicit QuietStderr(bool on) {
        if (!on) return;
        fflush(stderr);
        saved = dup(2);
        int nul = open("/dev/null", O_WRONLY);
        if (nul >= 0) { dup2(nul, 2); close(nul); }
    }
    ~QuietStderr() {
        if (saved < 0) return;
        fflush(stderr);
        dup2(saved, 2);
        close(saved);
    }
};

struct Cwd {
    std::string old;
    explicit Cwd(const char* dir) {
        if (!dir || !*dir) return;
        char buf[4096];
        if (getcwd(buf, sizeof(buf))) old = buf;
        if (chdir(dir) != 0) old.clear();
    }
    ~Cwd() { if (!old.empty() && chdir(old.c_str()) != 0) {} }
};

// ---- flattened copy of the last launch -------------------------------------------------
struct Flat {
    std::vector<rtw_prim> prims;
    std::vector<rtw_transform> xforms;
    std::vector<rtw_material> mats;
    std::vector<rtw_texture> texs;
    std::vector<rtw_perlin> perlin;
    std::vector<rtw_image> images;
    std::vector<std::vector<uint8_t>> image_data;
    std::vector<rtw_light> lights;
    std::vector<float> aabbs;           // 6 per instance, as handed to optixAccelBuild
    std::vector<int32_t> instance_ids;  // OptixInstance::instanceId per instance
    std::vector<int32_t> sbt_offsets;
    rtw_scene_desc desc{};
    int pdf_gen = -1, p1_gen = -1;
    std::string error;
};
Flat g_flat;

void put3(float* o, const float3& v) { o[0] = v.x; o[1] = v.y; o[2] = v.z; }

int prim_type_of(const std::string& is_name) {
    if (is_name == "__intersection__sphere") return RTW_PRIM_SPHERE;
    if (is_name == "__intersection__movingsphere") return RTW_PRIM_MOVING_SPHERE;
    if (is_name == "__intersection__hitRectX") return RTW_PRIM_RECT_X;
    if (is_name == "__intersection__hitRectY") return RTW_PRIM_RECT_Y;
    if (is_name == "__intersection__hitRectZ") return RTW_PRIM_RECT_Z;
    if (is_name == "__intersection__hitVolumebox") return RTW_PRIM_VOLUME_BOX;
    if (is_name == "__intersection__hitVolumesphere") return RTW_PRIM_VOLUME_SPHERE;
    return -1;
}
int material_type_of(int matindex) {   // shaders/FunctionIdx.h:17-25 -> lib/raydata.cuh:23-30
    switch (matindex) {
    case CALLABLE_ID_LAMBERTIAN: return RTW_MAT_LAMBERTIAN;
    case CALLABLE_ID_LIGHT_DIFFUSE: return RTW_MAT_DIFFUSE_LIGHT;
    case CALLABLE_ID_METAL: return RTW_MAT_METAL;
    case CALLABLE_ID_DIELECTRIC: return RTW_MAT_DIELECTRIC;
    case CALLABLE_ID_ISOTROPIC: return RTW_MAT_ISOTROPIC;
    case CALLABLE_ID_NORMAL: return RTW_MAT_NORMAL;
    }
    return -1;
}

// Flatten what the last optixLaunch ran on.  `Parameter` is the launch-parameter block the
// launch copied in; the SBT and the IAS are the ones Director::createSBT / ioScene built.
bool flatten_last_launch(Flat& f) {
    f = Flat();
    const OptixShaderBindingTable& sbt = refshim::last_sbt();
    const OptixTraversableHandle ias = Parameter.handle;
    const int n = refshim::ias_size(ias);
    if (n < 0) { f.error = "no IAS in the launch parameters"; return false; }
    std::map<std::string, int> tex_index;   // textures are per material in the reference; kept 1:1
    unsigned max_id = 0;
    for (int i = 0; i < n; i++) {
        const float* w2o; int identity;
        const OptixInstance* in = refshim::ias_instance(ias, i, &w2o, &identity);
        if (in->instanceId > max_id) max_id = in->instanceId;
        if (in->sbtOffset >= sbt.hitgroupRecordCount) { f.error = "instance " + std::to_string(i) + ": sbtOffset outside the SBT"; return false; }
        const char* rec = (const char*)sbt.hitgroupRecordBase + (size_t)in->sbtOffset * sbt.hitgroupRecordStrideInBytes;
        const HitGroupData* hg = (const HitGroupData*)(rec + OPTIX_SBT_RECORD_HEADER_SIZE);
        rtw_prim p{};
        p.type = prim_type_of(refshim::program_name(rec));
        if (p.type < 0) { f.error = "instance " + std::to_string(i) + ": unknown intersection program"; return false; }
        p.material = (int32_t)in->instanceId;     // Parameter.MatParams[optixGetInstanceId()], shaders/closehit.cu:49,64
        p.transform = -1;
        OptixAabb box{};
        const OptixMatrixMotionTransform* mt = nullptr;
        const int child = refshim::child_info(in->traversableHandle, &box, &mt);
        switch (p.type) {
        case RTW_PRIM_SPHERE:
            put3(p.v, hg->center); p.v[3] = hg->radius;
            break;
        case RTW_PRIM_MOVING_SPHERE:
            put3(p.v, hg->center); p.v[3] = hg->radius; put3(p.v + 4, hg->center1); p.v[7] = hg->time0; p.v[8] = hg->time1;
            if (child != 1 || !mt) { f.error = "moving sphere without a motion transform"; return false; }
            // the motion keys must be the translations c0 / c1 (geometry/ioMovingSphere.h:162-182, Q24)
            if (mt->transform[0][3] != hg->center.x || mt->transform[0][7] != hg->center.y || mt->transform[0][11] != hg->center.z ||
                mt->transform[1][3] != hg->center1.x || mt->transform[1][7] != hg->center1.y || mt->transform[1][11] != hg->center1.z) {
                f.error = "motion keys differ from the sphere centres"; return false;
            }
            break;
        case RTW_PRIM_RECT_X: case RTW_PRIM_RECT_Y: case RTW_PRIM_RECT_Z:
            p.v[0] = hg->rectData.a0; p.v[1] = hg->rectData.a1; p.v[2] = hg->rectData.b0; p.v[3] = hg->rectData.b1; p.v[4] = hg->rectData.k;
            p.flip = hg->rectData.flip;
            break;
        case RTW_PRIM_VOLUME_BOX:
            put3(p.v, hg->VolBoxData.boxMin); put3(p.v + 3, hg->VolBoxData.boxMax); p.v[6] = hg->VolBoxData.density;
            break;
        case RTW_PRIM_VOLUME_SPHERE:
            put3(p.v, hg->center); p.v[3] = hg->radius; p.v[4] = hg->density;
            break;
        }
        if (child == 1 && p.type != RTW_PRIM_MOVING_SPHERE) { f.error = "motion transform over a static primitive"; return false; }
        if (!identity) {
            rtw_transform t;
            memcpy(t.o2w, in->transform, sizeof(t.o2w));
            memcpy(t.w2o, w2o, sizeof(t.w2o));
            int found = -1;
            for (size_t k = 0; k < f.xforms.size(); k++)
                if (memcmp(&f.xforms[k], &t, sizeof(t)) == 0) { found = (int)k; break; }
            if (found < 0) { f.xforms.push_back(t); found = (int)f.xforms.size() - 1; }
            p.transform = found;
        }
        f.prims.push_back(p);
        const float b[6] = {box.minX, box.minY, box.minZ, box.maxX, box.maxY, box.maxZ};
        f.aabbs.insert(f.aabbs.end(), b, b + 6);
        f.instance_ids.push_back((int32_t)in->instanceId);
        f.sbt_offsets.push_back((int32_t)in->sbtOffset);
    }
    // materials, indexed by instanceId
    for (unsigned id = 0; id <= max_id; id++) {
        const MaterialParams& mp = Parameter.MatParams[id];
        rtw_material m{};
        m.type = material_type_of(mp.matindex);
        m.texture = -1;
        m.param = (m.type == RTW_MAT_METAL || m.type == RTW_MAT_DIELECTRIC) ? mp.fuzz : 0.f;
        if (m.type != RTW_MAT_DIELECTRIC && m.type != RTW_MAT_NORMAL && mp.texArr) {
            const textureParam& tp = mp.texArr[0];
            rtw_texture t{};
            t.type = tp.texCallidx;
            put3(t.color, tp.color);
            t.odd = t.even = t.perlin = t.image = -1;
            if (t.type == RTW_TEX_CHECKER) { t.odd = tp.odd; t.even = tp.even; }   // raw callable indices (Q18)
            if (t.type == RTW_TEX_NOISE) {
                t.scale = tp.scale;
                rtw_perlin pl;
                memcpy(pl.ranvec, (const void*)tp.randvec, sizeof(pl.ranvec));
                memcpy(pl.perm_x, (const void*)tp.permX, sizeof(pl.perm_x));
                memcpy(pl.perm_y, (const void*)tp.permY, sizeof(pl.perm_y));
                memcpy(pl.perm_z, (const void*)tp.permZ, sizeof(pl.perm_z));
                f.perlin.push_back(pl);
                t.perlin = (int)f.perlin.size() - 1;
            }
            if (t.type == RTW_TEX_IMAGE) {
                const cudaArray* a = refshim::texture_array(tp.imagetex);
                rtw_image im{};
                im.width = a->width; im.height = a->height;
                f.image_data.emplace_back(a->rgba, a->rgba + (size_t)a->width * a->height * 4);
                f.images.push_back(im);
                t.image = (int)f.images.size() - 1;
            }
            f.texs.push_back(t);
            m.texture = (int)f.texs.size() - 1;
        }
        f.mats.push_back(m);
    }
    for (size_t i = 0; i < f.images.size(); i++) f.images[i].rgba = f.image_data[i].data();
    for (int i = 0; i < Parameter.numLights; i++) {
        const LightDefinition& L = Parameter.lights[i];
        rtw_light l{};
        put3(l.position, L.position); put3(l.vecU, L.vecU); put3(l.vecV, L.vecV); put3(l.normal, L.normal);
        l.area = L.area; put3(l.emission, L.emission);
        f.lights.push_back(l);
    }
    rtw_scene_desc& d = f.desc;
    d.n_prims = (int)f.prims.size(); d.prims = f.prims.data();
    d.n_transforms = (int)f.xforms.size(); d.transforms = f.xforms.data();
    d.n_materials = (int)f.mats.size(); d.materials = f.mats.data();
    d.n_textures = (int)f.texs.size(); d.textures = f.texs.data();
    d.n_perlin = (int)f.perlin.size(); d.perlin = f.perlin.data();
    d.n_images = (int)f.images.size(); d.images = f.images.data();
    d.n_lights = (int)f.lights.size(); d.lights = f.lights.data();
    d.sky_light = Parameter.skyLight;
    f.pdf_gen = Parameter.pdf.pdfGenIdx;
    d.has_light_rect = 0;
    if (Parameter.numLights > 0 && Parameter.pdf.pdfGenIdx == CALLABLE_ID_mixture_generate && Parameter.pdf.p1) {
        // __direct_callable__mixture_generate forwards to p1 with p1's rectangle (pdf/mixturePdf.cu:25-38)
        const pdfCallfun* p1 = Parameter.pdf.p1;
        f.p1_gen = p1->pdfGenIdx;
        d.has_light_rect = 1;
        d.light_rect.axis = p1->pdfGenIdx == CALLABLE_ID_rect_X_generate ? 0 : p1->pdfGenIdx == CALLABLE_ID_rect_Y_generate ? 1 : 2;
        d.light_rect.a0 = p1->pdfrect.a0; d.light_rect.a1 = p1->pdfrect.a1;
        d.light_rect.b0 = p1->pdfrect.b0; d.light_rect.b1 = p1->pdfrect.b1; d.light_rect.k = p1->pdfrect.k;
    }
    rtw_camera& c = d.camera;
    put3(c.origin, Parameter.cameraOrigin); put3(c.u, Parameter.cameraU); put3(c.v, Parameter.cameraV); put3(c.w, Parameter.cameraW);
    put3(c.lower_left, Parameter.cameraLowerLeftCorner); put3(c.horizontal, Parameter.cameraHorizontal); put3(c.vertical, Parameter.cameraVertical);
    c.lens_radius = Parameter.cameraLensRadius; c.time0 = Parameter.cameraTime0; c.time1 = Parameter.cameraTime1;
    return true;
}

// ---- building a launch from an rtw_scene_desc (ref_render_desc) -----------------------
struct HostRecordHeader { __align__(OPTIX_SBT_RECORD_ALIGNMENT) char header[OPTIX_SBT_RECORD_HEADER_SIZE]; };
struct HostHitRecord { __align__(OPTIX_SBT_RECORD_ALIGNMENT) char header[OPTIX_SBT_RECORD_HEADER_SIZE]; HitGroupData data; };

const char* kIsNames[RTW_PRIM_TYPE_COUNT] = {"__intersection__sphere", "__intersection__movingsphere", "__intersection__hitRectX",
                                             "__intersection__hitRectY", "__intersection__hitRectZ", "__intersection__hitVolumebox",
                                             "__intersection__hitVolumesphere"};
// callables in SBT order: camera, materials (MatCallFunction), textures (TexCallFunction), pdfs (PDFCallFunction)
const char* kCallableNames[] = {
    "__direct_callable__perspectiveCamera",
    "__direct_callable__diffuselight", "__direct_callable__dielectricMat", "__direct_callable__ISOTROPIC", "__direct_callable__lambertian",
    "__direct_callable__metal", "__direct_callable__normal",
    "__direct_callable__checkerTexture", "__direct_callable__constantTexture", "__direct_callable__imagetexture", "__direct_callable__noisetexture",
    "__direct_callable__nullTexture",
    "__direct_callable__cosineGenerate", "__direct_callable__mixtureBIAS_generate", "__direct_callable__mixture_generate",
    "__direct_callable__rect_x_generate", "__direct_callable__rect_y_generate", "__direct_callable__rect_z_generate",
    "__direct_callable__cosineValue", "__direct_callable__mixtureBIAS_value", "__direct_callable__mixture_value",
    "__direct_callable__rect_x_value", "__direct_callable__rect_y_value", "__direct_callable__rect_z_value",
    "__direct_callable__eval_bsdf_diffuse_reflection", "__direct_callable__eval_bsdf_dielect_reflection",
    "__direct_callable__eval_bsdf_metal_reflection"};

}  // namespace

int ref_main(int argc, char* argv[]);   // the reference's main(), RestOfLife/main.cpp:29 (built with -Dmain=ref_main)

extern "C" {

int ref_run_main(int argc, char** argv) { return ref_main(argc, argv); }

int ref_set_threads(int n) { return refshim::set_threads(n); }
int ref_set_q23_normalise(int on) { return refshim::set_q23_normalise(on); }
unsigned long long ref_traversal_count() { return refshim::traversal_count(); }
const char* ref_last_error() { return g_flat.error.c_str(); }

/* Director::init(Nx, Ny, 1) -> createScene(scene) -> renderFrame() `launches` times (each launch is
 * the same 1 spp frame: the raygen program seeds tea<64>(pixel, 0), raygen/raygen.cu:129).
 * asset_root: directory that holds assets/earthmap.jpg (the reference opens a relative path,
 * scene/ioScene.h:357,795).  rgba_out: Nx*Ny*4 floats, the launch's frame_buffer. */
int ref_render_scene(int scene, int Nx, int Ny, int launches, const char* asset_root, int quiet, float* rgba_out) {
    try {
        Cwd cwd(asset_root);
        Director* d = new Director(false, false);   // not destroyed: Director::destroy double-frees (Director.cpp:66-104)
        {
            QuietStderr q(quiet != 0);
            d->init(Nx, Ny, 1);
            d->createScene(scene);
        }
        for (int i = 0; i < (launches > 0 ? launches : 1); i++) d->renderFrame();
        if (rgba_out) memcpy(rgba_out, Parameter.frame_buffer, sizeof(float4) * (size_t)Nx * Ny);
        if (!flatten_last_launch(g_flat)) return 2;
        return 0;
    } catch (const std::exception& e) {
        g_flat.error = e.what();
        return 1;
    }
}

/* A session for timing (bench.py's reference arm): Director::init + createScene once, then renderFrame()
 * as often as asked.  Every launch renders the same 1 spp frame (raygen/raygen.cu:129 seeds sample 0). */
static Director* g_session = nullptr;
static int g_session_nx = 0, g_session_ny = 0;
int ref_open(int scene, int Nx, int Ny, const char* asset_root, int quiet) {
    try {
        Cwd cwd(asset_root);
        QuietStderr q(quiet != 0);
        g_session = new Director(false, false);
        g_session->init(Nx, Ny, 1);
        g_session->createScene(scene);
        g_session_nx = Nx; g_session_ny = Ny;
        return 0;
    } catch (const std::exception& e) {
        g_flat.error = e.what();
        g_session = nullptr;
        return 1;
    }
}
int ref_launch(int launches) {
    if (!g_session) return 1;
    try {
        for (int i = 0; i < launches; i++) g_session->renderFrame();
        return 0;
    } catch (const std::exception& e) {
        g_flat.error = e.what();
        return 1;
    }
}
int ref_frame(float* rgba_out) {
    if (!g_session || !rgba_out) return 1;
    memcpy(rgba_out, Parameter.frame_buffer, sizeof(float4) * (size_t)g_session_nx * g_session_ny);
    return 0;
}

/* The scene the last ref_render_scene launch ran on, as an rtw_scene_desc (owned by the library,
 * valid until the next call).  prims[i] = instance i of the IAS; prims[i].material = its instanceId. */
const rtw_scene_desc* ref_scene_desc() { return g_flat.prims.empty() ? nullptr : &g_flat.desc; }
/* extras that rtw_scene_desc has no field for */
int ref_scene_extras(float* aabbs6, int32_t* instance_ids, int32_t* sbt_offsets, int32_t* pdf_gen, int32_t* p1_gen) {
    if (aabbs6) memcpy(aabbs6, g_flat.aabbs.data(), g_flat.aabbs.size() * sizeof(float));
    if (instance_ids) memcpy(instance_ids, g_flat.instance_ids.data(), g_flat.instance_ids.size() * sizeof(int32_t));
    if (sbt_offsets) memcpy(sbt_offsets, g_flat.sbt_offsets.data(), g_flat.sbt_offsets.size() * sizeof(int32_t));
    if (pdf_gen) *pdf_gen = g_flat.pdf_gen;
    if (p1_gen) *p1_gen = g_flat.p1_gen;
    return (int)g_flat.prims.size();
}

/* The reference's device programs on an arbitrary scene description.  One launch = 1 spp with
 * the seeds of sample 0.  Checker textures: `odd` / `even` must name a CONSTANT or NULL texture;
 * the reference's checker hands its own parameter block to the child (texture/checkeredTexture.cu:
 * 15-18 passes matPar, whose texArr[0] is the checker), so a constant child reads the checker's
 * `color` field — at most one distinct constant colour per checker can be expressed, and the
 * child callable index is made absolute here (Q18). */
int ref_render_desc(const rtw_scene_desc* s, int Nx, int Ny, int max_depth, float* rgba_out) {
    try {
        OptixDeviceContext ctx;
        optixDeviceContextCreate(nullptr, nullptr, &ctx);
        // ---- program groups ----
        const int n_call = (int)(sizeof(kCallableNames) / sizeof(kCallableNames[0]));
        if (n_call != NUM_CALLABLE_CAMERA + NUM_CALLABLE_MAT_IDS + NUM_CALLABE_TEX_IDS + NUM_CALLABLE_PDF_IDS) { g_flat.error = "callable table"; return 1; }
        std::vector<OptixProgramGroupDesc> descs(2 + RTW_PRIM_TYPE_COUNT + n_call);
        memset(descs.data(), 0, descs.size() * sizeof(OptixProgramGroupDesc));
        descs[0].kind = OPTIX_PROGRAM_GROUP_KIND_RAYGEN; descs[0].raygen.entryFunctionName = "__raygen__Program";
        descs[1].kind = OPTIX_PROGRAM_GROUP_KIND_MISS; descs[1].miss.entryFunctionName = "__miss__Program";
        for (int t = 0; t < RTW_PRIM_TYPE_COUNT; t++) {
            descs[2 + t].kind = OPTIX_PROGRAM_GROUP_KIND_HITGROUP;
            descs[2 + t].hitgroup.entryFunctionNameCH = "__closesthit__radiance";
            descs[2 + t].hitgroup.entryFunctionNameIS = kIsNames[t];
        }
        for (int c = 0; c < n_call; c++) {
            descs[2 + RTW_PRIM_TYPE_COUNT + c].kind = OPTIX_PROGRAM_GROUP_KIND_CALLABLES;
            descs[2 + RTW_PRIM_TYPE_COUNT + c].callables.entryFunctionNameDC = kCallableNames[c];
        }
        std::vector<OptixProgramGroup> pgs(descs.size());
        optixProgramGroupCreate(ctx, descs.data(), (unsigned)descs.size(), nullptr, nullptr, nullptr, pgs.data());
        // ---- SBT ----
        static std::vector<HostRecordHeader> rg(1), ms(1), dc;
        static std::vector<HostHitRecord> hit;
        optixSbtRecordPackHeader(pgs[0], &rg[0]);
        optixSbtRecordPackHeader(pgs[1], &ms[0]);
        dc.assign(n_call, HostRecordHeader{});
        for (int c = 0; c < n_call; c++) optixSbtRecordPackHeader(pgs[2 + RTW_PRIM_TYPE_COUNT + c], &dc[c]);
        hit.assign(s->n_prims, HostHitRecord{});
        // ---- geometry: one GAS + one instance per primitive, as the reference does ----
        static std::vector<OptixInstance> inst;
        static std::vector<OptixMatrixMotionTransform> motions;
        inst.assign(s->n_prims, OptixInstance{});
        motions.assign(s->n_prims, OptixMatrixMotionTransform{});
        for (int i = 0; i < s->n_prims; i++) {
            const rtw_prim& p = s->prims[i];
            if (p.type < 0 || p.type >= RTW_PRIM_TYPE_COUNT) { g_flat.error = "bad primitive type"; return 1; }
            optixSbtRecordPackHeader(pgs[2 + p.type], &hit[i]);
            HitGroupData& hg = hit[i].data;
            memset(&hg, 0, sizeof(hg));
            switch (p.type) {
            case RTW_PRIM_SPHERE: hg.center = make_float3(p.v[0], p.v[1], p.v[2]); hg.radius = p.v[3]; break;
            case RTW_PRIM_MOVING_SPHERE:
                hg.center = make_float3(p.v[0], p.v[1], p.v[2]); hg.radius = p.v[3]; hg.center1 = make_float3(p.v[4], p.v[5], p.v[6]);
                hg.time0 = p.v[7]; hg.time1 = p.v[8]; hg.bMotion = true;
                break;
            case RTW_PRIM_RECT_X: case RTW_PRIM_RECT_Y: case RTW_PRIM_RECT_Z:
                hg.rectData.a0 = p.v[0]; hg.rectData.a1 = p.v[1]; hg.rectData.b0 = p.v[2]; hg.rectData.b1 = p.v[3]; hg.rectData.k = p.v[4];
                hg.rectData.flip = p.flip;
                break;
            case RTW_PRIM_VOLUME_BOX:
                hg.VolBoxData.boxMin = make_float3(p.v[0], p.v[1], p.v[2]); hg.VolBoxData.boxMax = make_float3(p.v[3], p.v[4], p.v[5]);
                hg.VolBoxData.density = p.v[6];
                break;
            case RTW_PRIM_VOLUME_SPHERE: hg.center = make_float3(p.v[0], p.v[1], p.v[2]); hg.radius = p.v[3]; hg.density = p.v[4]; break;
            }
            OptixAabb box{-1e30f, -1e30f, -1e30f, 1e30f, 1e30f, 1e30f};   // the stand-in does not cull
            CUdeviceptr boxp = (CUdeviceptr)&box;
            OptixBuildInput bi{};
            bi.type = OPTIX_BUILD_INPUT_TYPE_CUSTOM_PRIMITIVES;
            bi.customPrimitiveArray.aabbBuffers = &boxp;
            bi.customPrimitiveArray.numPrimitives = 1;
            OptixTraversableHandle gas;
            optixAccelBuild(ctx, nullptr, nullptr, &bi, 1, 0, 0, 0, 0, &gas, nullptr, 0);
            OptixInstance& in = inst[i];
            const float ident[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
            memcpy(in.transform, p.transform >= 0 ? s->transforms[p.transform].o2w : ident, sizeof(ident));
            in.instanceId = (unsigned)i;
            in.sbtOffset = (unsigned)i;
            in.visibilityMask = 255;
            in.traversableHandle = gas;
            if (p.type == RTW_PRIM_MOVING_SPHERE) {   // geometry/ioMovingSphere.h:162-200
                OptixMatrixMotionTransform& mt = motions[i];
                mt.child = gas;
                mt.motionOptions.numKeys = 2; mt.motionOptions.timeBegin = 0.f; mt.motionOptions.timeEnd = 1.f;
                const float k0[12] = {1, 0, 0, p.v[0], 0, 1, 0, p.v[1], 0, 0, 1, p.v[2]};
                const float k1[12] = {1, 0, 0, p.v[4], 0, 1, 0, p.v[5], 0, 0, 1, p.v[6]};
                memcpy(mt.transform[0], k0, sizeof(k0)); memcpy(mt.transform[1], k1, sizeof(k1));
                optixConvertPointerToTraversableHandle(ctx, (CUdeviceptr)&mt, OPTIX_TRAVERSABLE_TYPE_MATRIX_MOTION_TRANSFORM, &in.traversableHandle);
            }
        }
        OptixBuildInput ii{};
        ii.type = OPTIX_BUILD_INPUT_TYPE_INSTANCES;
        ii.instanceArray.instances = (CUdeviceptr)inst.data();
        ii.instanceArray.numInstances = (unsigned)inst.size();
        OptixTraversableHandle ias;
        optixAccelBuild(ctx, nullptr, nullptr, &ii, 1, 0, 0, 0, 0, &ias, nullptr, 0);
        // ---- launch parameters (Director::init + initLaunchParams, Director.cpp:33-48,483-553) ----
        static std::vector<MaterialParams> mats;
        static std::vector<textureParam> texs;
        static std::vector<LightDefinition> lights;
        static pdfCallfun p0, p1;
        static std::vector<float4> frame;
        const int tex_base = NUM_CALLABLE_CAMERA + NUM_CALLABLE_MAT_IDS;
        texs.assign(s->n_prims, textureParam{});
        mats.assign(s->n_prims, MaterialParams{});
        for (int i = 0; i < s->n_prims; i++) {   // MatParams is indexed by instanceId == primitive index
            const rtw_material& m = s->materials[s->prims[i].material];
            MaterialParams& mp = mats[i];
            mp.texArr = nullptr;
            switch (m.type) {   // material/io*Material.h assignTo
            case RTW_MAT_LAMBERTIAN: mp.matindex = CALLABLE_ID_LAMBERTIAN; mp.lightreflectIdx = CALLABLE_ID_LIGHT_SAMPLE_PDF; break;
            case RTW_MAT_DIFFUSE_LIGHT: mp.matindex = CALLABLE_ID_LIGHT_DIFFUSE; break;
            case RTW_MAT_METAL: mp.matindex = CALLABLE_ID_METAL; mp.lightreflectIdx = CALLABLE_ID_LIGHT_SAMPLE_METAL_PDF; mp.fuzz = m.param; break;
            case RTW_MAT_DIELECTRIC: mp.matindex = CALLABLE_ID_DIELECTRIC; mp.lightreflectIdx = CALLABLE_ID_LIGHT_SAMPLE_DIELECT_PDF; mp.eta = m.param; break;
            case RTW_MAT_ISOTROPIC: mp.matindex = CALLABLE_ID_ISOTROPIC; break;
            default: mp.matindex = CALLABLE_ID_NORMAL; break;
            }
            if (m.texture >= 0) {
                const rtw_texture& t = s->textures[m.texture];
                textureParam& tp = texs[i];
                tp.texCallidx = t.type;
                tp.color = make_float3(t.color[0], t.color[1], t.color[2]);
                tp.scale = t.scale;
                if (t.type == RTW_TEX_CHECKER) {
                    const rtw_texture& o = s->textures[t.odd];
                    const rtw_texture& e = s->textures[t.even];
                    if ((o.type != RTW_TEX_CONSTANT && o.type != RTW_TEX_NULL) || (e.type != RTW_TEX_CONSTANT && e.type != RTW_TEX_NULL)) {
                        g_flat.error = "checker children must be constant or null textures"; return 1;
                    }
                    const rtw_texture* cst = o.type == RTW_TEX_CONSTANT ? &o : (e.type == RTW_TEX_CONSTANT ? &e : nullptr);
                    if (o.type == RTW_TEX_CONSTANT && e.type == RTW_TEX_CONSTANT && memcmp(o.color, e.color, sizeof(o.color)) != 0) {
                        g_flat.error = "the reference's checker can express one constant colour only"; return 1;
                    }
                    if (cst) tp.color = make_float3(cst->color[0], cst->color[1], cst->color[2]);
                    tp.odd = o.type + tex_base;
                    tp.even = e.type + tex_base;
                }
                if (t.type == RTW_TEX_NOISE) {
                    const rtw_perlin& pl = s->perlin[t.perlin];
                    tp.randvec = (CUdeviceptr)pl.ranvec; tp.permX = (CUdeviceptr)pl.perm_x; tp.permY = (CUdeviceptr)pl.perm_y; tp.permZ = (CUdeviceptr)pl.perm_z;
                }
                if (t.type == RTW_TEX_IMAGE) {
                    const rtw_image& im = s->images[t.image];
                    cudaArray_t arr;
                    cudaMallocArray(&arr, nullptr, im.width, im.height);
                    cudaMemcpy2DToArray(arr, 0, 0, im.rgba, (size_t)im.width * 4, (size_t)im.width * 4, im.height, cudaMemcpyHostToDevice);
                    cudaResourceDesc rd{};
                    rd.res.array.array = arr;
                    cudaCreateTextureObject(&tp.imagetex, &rd, nullptr, nullptr);
                }
                mp.texArr = &texs[i];
            }
        }
        SysParamter P{};
        P.handle = ias;
        P.m_Nx = Nx; P.m_Ny = Ny;
        frame.assign((size_t)Nx * Ny, float4{});
        P.frame_buffer = frame.data();
        P.maxRayDepth = max_depth;
        P.numSamples = 1;
        P.cameraType = CALLABLE_ID_PINHOLE;
        lights.clear();
        for (int i = 0; i < s->n_lights; i++) {
            const rtw_light& l = s->lights[i];
            LightDefinition L{};
            L.position = make_float3(l.position[0], l.position[1], l.position[2]);
            L.vecU = make_float3(l.vecU[0], l.vecU[1], l.vecU[2]); L.vecV = make_float3(l.vecV[0], l.vecV[1], l.vecV[2]);
            L.normal = make_float3(l.normal[0], l.normal[1], l.normal[2]);
            L.area = l.area; L.emission = make_float3(l.emission[0], l.emission[1], l.emission[2]);
            lights.push_back(L);
        }
        P.lights = lights.data();
        P.numLights = s->has_light_rect ? (int)lights.size() : 0;   // the oracle samples lights only through the rectangle
        P.skyLight = s->sky_light;
        if (P.numLights > 0) {   // scene/ioScene.h:103-148: mixture(cosine, rect)
            P.pdf = pdfCallfun{};
            P.pdf.pdfGenIdx = CALLABLE_ID_mixture_generate; P.pdf.pdfValIdx = CALLABLE_ID_mixture_value;
            p0 = pdfCallfun{}; p0.pdfGenIdx = CALLABLE_ID_cosineGenerate; p0.pdfValIdx = CALLABLE_ID_cosineValue;
            p1 = pdfCallfun{};
            const int ax = s->light_rect.axis;
            p1.pdfGenIdx = ax == 0 ? CALLABLE_ID_rect_X_generate : ax == 1 ? CALLABLE_ID_rect_Y_generate : CALLABLE_ID_rect_Z_generate;
            p1.pdfValIdx = ax == 0 ? CALLABLE_ID_rect_x_value : ax == 1 ? CALLABLE_ID_rect_y_value : CALLABLE_ID_rect_z_value;
            p1.pdfrect = hitRectData{s->light_rect.a0, s->light_rect.a1, s->light_rect.b0, s->light_rect.b1, s->light_rect.k, 0};
            P.pdf.p0 = &p0; P.pdf.p1 = &p1;
            P.pdf.pdfrect = p1.pdfrect;
        }
        const rtw_camera& c = s->camera;
        P.cameraOrigin = make_float3(c.origin[0], c.origin[1], c.origin[2]);
        P.cameraU = make_float3(c.u[0], c.u[1], c.u[2]); P.cameraV = make_float3(c.v[0], c.v[1], c.v[2]); P.cameraW = make_float3(c.w[0], c.w[1], c.w[2]);
        P.cameraTime0 = c.time0; P.cameraTime1 = c.time1; P.cameraLensRadius = c.lens_radius;
        P.cameraLowerLeftCorner = make_float3(c.lower_left[0], c.lower_left[1], c.lower_left[2]);
        P.cameraHorizontal = make_float3(c.horizontal[0], c.horizontal[1], c.horizontal[2]);
        P.cameraVertical = make_float3(c.vertical[0], c.vertical[1], c.vertical[2]);
        P.MatParams = mats.data();
        OptixShaderBindingTable sbt{};
        sbt.raygenRecord = (CUdeviceptr)rg.data();
        sbt.missRecordBase = (CUdeviceptr)ms.data(); sbt.missRecordStrideInBytes = sizeof(HostRecordHeader); sbt.missRecordCount = 1;
        sbt.hitgroupRecordBase = (CUdeviceptr)hit.data(); sbt.hitgroupRecordStrideInBytes = sizeof(HostHitRecord); sbt.hitgroupRecordCount = (unsigned)hit.size();
        sbt.callablesRecordBase = (CUdeviceptr)dc.data(); sbt.callablesRecordStrideInBytes = sizeof(HostRecordHeader); sbt.callablesRecordCount = (unsigned)n_call;
        optixLaunch(nullptr, nullptr, (CUdeviceptr)&P, sizeof(P), &sbt, (unsigned)Nx, (unsigned)Ny, 1);
        if (rgba_out) memcpy(rgba_out, frame.data(), sizeof(float4) * frame.size());
        return 0;
    } catch (const std::exception& e) {
        g_flat.error = e.what();
        return 1;
    }
}

/* RestOfLife/texture/ioTexture.h:228-266 on its own: decode + flip + RGBA8 through the reference's
 * ioImageTexture (external/stb_image.h).  Returns the texel array of the texture object it creates. */
int ref_load_image(const char* asset_root, const char* file, int* w, int* h, unsigned char* rgba_out, int capacity) {
    try {
        Cwd cwd(asset_root);
        QuietStderr q(true);
        ioImageTexture* t = new ioImageTexture(file);
        const cudaArray* a = refshim::texture_array(t->TexObj);
        *w = a->width; *h = a->height;
        if (rgba_out) {
            if (capacity < a->width * a->height * 4) return 2;
            memcpy(rgba_out, a->rgba, (size_t)a->width * a->height * 4);
        }
        return 0;
    } catch (const std::exception& e) {
        g_flat.error = e.what();
        return 1;
    }
}

}  // extern "C"
