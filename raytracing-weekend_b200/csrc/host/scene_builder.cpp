// scene_builder.cpp — host-side scene flattening behind rtw_scene_create().
//
// Restates the reference's scene functions as flat POD arrays (include/rtw.h):
//   scene/ioScene.h:52-101   ioScene::init dispatch          → rtw_scene_create
//   scene/ioScene.h:103-148  getScenePdf                     → light_rect
//   scene/ioScene.h:180-304  MovingSpheres        (scene 1)
//   scene/ioScene.h:308-493  InOneWeekendLight    (scene 2)
//   scene/ioScene.h:495-629  CornellBox           (scene 0)
//   scene/ioScene.h:632-788  VolumesCornellBox    (scene 3)
//   scene/ioScene.h:791-982  TheNextWeekFinal     (scene 4)
//   scene/ioCamera.h:64-90   ioPerspectiveCamera
//   geometry/ioTransform.h:68-131, geometry/ioGeometryGroup.h:27-40 (createBox)
//   texture/ioTexture.h:21-26,128-216 (Perlin tables), :228-266 (image flip)
//   Director.cpp:483-553     initLaunchParams (camera times, skyLight, lights)
// plus the synthetic stress scene C5 of SURVEY.md §8d (scene 5).
//
// No OptiX handles: one flat primitive array {type, params, material, transform}
// replaces the per-primitive GAS + instance + SBT-record bookkeeping.
// Primitive order == the reference's instanceId order, so prim index i is the
// reference's MatParams index i.
//
// Documented normalisations: Q33 argument-evaluation order fixed to
// right-to-left (the MSVC x64 behaviour the reference was built with); Q19 Perlin
// tables from std::mt19937(0) raw outputs / 2^32 (libstdc++'s
// generate_canonical<float,24> result), vector components drawn right-to-left.
#include "../../../include/rtw.h"

#include <cmath>
#include <cstdio>
#include <cstring>
#include <random>
#include <string>
#include <vector>

namespace {

// ---- minimal float3 helpers (sutil/vec_math.h semantics: v/s = v*(1/s)) ----
struct F3 { float x, y, z; };
inline F3 f3(float x, float y, float z) { return F3{x, y, z}; }
inline F3 operator+(F3 a, F3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline F3 operator-(F3 a, F3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline F3 operator*(float s, F3 a) { return f3(s * a.x, s * a.y, s * a.z); }
inline F3 operator*(F3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
inline float dot(F3 a, F3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline F3 cross(F3 a, F3 b) { return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline float length(F3 a) { return sqrtf(dot(a, a)); }
inline F3 normalize(F3 a) { float inv = 1.0f / sqrtf(dot(a, a)); return a * inv; }
inline void put(float* d, F3 v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; }

const float kPi = 3.141592654f;  // CUDART_PI_F

// ---- lib/random.cuh:22-38 (host use: scene/ioScene.h:33) ----
inline uint32_t xorshift32(uint32_t& s) {
    s ^= s << 13;
    s ^= s >> 17;
    s ^= s << 5;
    return s;
}
inline float randf(uint32_t& s) {
    float r = ((float)xorshift32(s)) / 4294967296.0f;
    if (r != 1.0f) return r;
    return (float)0x3F7FFFFF;  // literal: an int constant converted to float
}

// ---- sutil::Matrix4x4 (row-major; operator*= is m1 = m1 * m2) ----
struct M4 { float m[16]; };
inline M4 mul(const M4& a, const M4& b) {
    M4 r;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            float sum = 0.0f;
            for (int k = 0; k < 4; k++) sum += a.m[i * 4 + k] * b.m[k * 4 + j];
            r.m[i * 4 + j] = sum;
        }
    return r;
}
// geometry/ioTransform.h:68-78
inline M4 translate(F3 o) {
    return M4{{1.0f, 0.0f, 0.0f, o.x, 0.0f, 1.0f, 0.0f, o.y, 0.0f, 0.0f, 1.0f, o.z, 0.f, 0.f, 0.f, 1.f}};
}
// geometry/ioTransform.h:28-32,105-116
inline M4 rotateY(float deg) {
    float a = deg * kPi / 180.f;
    return M4{{cosf(a), 0.0f, sinf(a), 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, -sinf(a), 0.0f, cosf(a), 0.0f, 0.0f, 0.0f, 0.0f, 1.0f}};
}
// geometry/ioTransform.h:41-46,119-129
inline M4 rotateZ(float deg) {
    float a = deg * kPi / 180.f;
    return M4{{cosf(a), -sinf(a), 0.0f, 0.0f, sinf(a), cosf(a), 0.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.f, 0.f, 0.f, 1.f}};
}

// OptiX inverts instance transforms internally; here the inverse is formed in
// double by cofactors and rounded once to float (DESIGN.md "transforms").
rtw_transform make_transform(const M4& m) {
    rtw_transform t;
    for (int i = 0; i < 12; i++) t.o2w[i] = m.m[i];
    double a[3][3], tr[3];
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) a[i][j] = m.m[i * 4 + j];
        tr[i] = m.m[i * 4 + 3];
    }
    double c00 = a[1][1] * a[2][2] - a[1][2] * a[2][1];
    double c01 = a[1][2] * a[2][0] - a[1][0] * a[2][2];
    double c02 = a[1][0] * a[2][1] - a[1][1] * a[2][0];
    double det = a[0][0] * c00 + a[0][1] * c01 + a[0][2] * c02;
    double inv[3][3];
    inv[0][0] = c00 / det;
    inv[1][0] = c01 / det;
    inv[2][0] = c02 / det;
    inv[0][1] = (a[0][2] * a[2][1] - a[0][1] * a[2][2]) / det;
    inv[1][1] = (a[0][0] * a[2][2] - a[0][2] * a[2][0]) / det;
    inv[2][1] = (a[0][1] * a[2][0] - a[0][0] * a[2][1]) / det;
    inv[0][2] = (a[0][1] * a[1][2] - a[0][2] * a[1][1]) / det;
    inv[1][2] = (a[0][2] * a[1][0] - a[0][0] * a[1][2]) / det;
    inv[2][2] = (a[0][0] * a[1][1] - a[0][1] * a[1][0]) / det;
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) t.w2o[i * 4 + j] = (float)inv[i][j];
        t.w2o[i * 4 + 3] = (float)(-(inv[i][0] * tr[0] + inv[i][1] * tr[1] + inv[i][2] * tr[2]));
    }
    return t;
}

// ---- texture/ioTexture.h:21-26: static mt19937(0) + uniform_real_distribution<float>(0,1) ----
struct LocalRnd {
    std::mt19937 gen{0};
    float operator()() {
        float r = (float)gen() / 4294967296.0f;
        if (r >= 1.0f) r = 0.99999994f;  // generate_canonical never returns 1
        return r;
    }
};

}  // namespace

struct rtw_scene {
    std::vector<rtw_prim> prims;
    std::vector<rtw_transform> xforms;
    std::vector<rtw_material> mats;
    std::vector<rtw_texture> texs;
    std::vector<rtw_perlin> perlin;
    std::vector<rtw_image> images;
    std::vector<std::vector<uint8_t>> image_data;
    std::vector<rtw_light> lights;
    rtw_scene_desc desc{};
    std::string description;
    std::string asset_dir;
    LocalRnd local_rnd;
    std::string error;

    // ---- textures ----
    int tex_constant(F3 c) {
        rtw_texture t{};
        t.type = RTW_TEX_CONSTANT;
        put(t.color, c);
        t.odd = t.even = t.perlin = t.image = -1;
        texs.push_back(t);
        return (int)texs.size() - 1;
    }
    int tex_checker(int odd, int even) {
        rtw_texture t{};
        t.type = RTW_TEX_CHECKER;
        t.odd = odd;
        t.even = even;
        t.perlin = t.image = -1;
        texs.push_back(t);
        return (int)texs.size() - 1;
    }
    // texture/ioTexture.h:181-216 (tables are drawn when the material is flattened)
    int tex_noise(float scale) {
        rtw_perlin p;
        for (int i = 0; i < 256; ++i) {
            // unit_float3(-1 + 2*localRnd(), ...): arguments evaluated right-to-left (Q19/Q33)
            float z = -1 + 2 * local_rnd();
            float y = -1 + 2 * local_rnd();
            float x = -1 + 2 * local_rnd();
            float l = sqrtf(x * x + y * y + z * z);
            p.ranvec[i][0] = x / l;
            p.ranvec[i][1] = y / l;
            p.ranvec[i][2] = z / l;
        }
        int32_t* perms[3] = {p.perm_x, p.perm_y, p.perm_z};
        for (int a = 0; a < 3; a++) {
            int32_t* pm = perms[a];
            for (int i = 0; i < 256; i++) pm[i] = i;
            for (int i = 256 - 1; i > 0; i--) {  // ioTexture.h:137-144
                int target = int(local_rnd() * (i + 1));
                int tmp = pm[i];
                pm[i] = pm[target];
                pm[target] = tmp;
            }
        }
        perlin.push_back(p);
        rtw_texture t{};
        t.type = RTW_TEX_NOISE;
        t.scale = scale;
        t.perlin = (int)perlin.size() - 1;
        t.odd = t.even = t.image = -1;
        texs.push_back(t);
        return (int)texs.size() - 1;
    }
    // texture/ioTexture.h:228-266: RGBA8, vertical flip; texels come from the
    // pre-decoded assets/earthmap.ppm (tools/make_earthmap_ppm.py).
    int tex_image(const std::string& file) {
        std::string path = asset_dir + "/" + file;
        FILE* f = fopen(path.c_str(), "rb");
        if (!f) { error = "cannot open " + path; return -1; }
        int w = 0, h = 0, maxv = 0;
        char magic[3] = {0, 0, 0};
        if (fscanf(f, "%2s %d %d %d", magic, &w, &h, &maxv) != 4 || strcmp(magic, "P6") != 0 || maxv != 255 || w <= 0 || h <= 0) {
            fclose(f);
            error = "not a binary PPM (P6, maxval 255): " + path;
            return -1;
        }
        fgetc(f);  // single whitespace after the header
        std::vector<uint8_t> rgb((size_t)w * h * 3);
        size_t got = fread(rgb.data(), 1, rgb.size(), f);
        fclose(f);
        if (got != rgb.size()) { error = "short read: " + path; return -1; }
        std::vector<uint8_t> rgba((size_t)w * h * 4);
        for (int i = 0; i < w; ++i)
            for (int j = 0; j < h; ++j) {
                size_t bindex = ((size_t)j * w + i);
                size_t iindex = ((size_t)(h - j - 1) * w + i) * 3;
                rgba[bindex * 4 + 0] = rgb[iindex + 0];
                rgba[bindex * 4 + 1] = rgb[iindex + 1];
                rgba[bindex * 4 + 2] = rgb[iindex + 2];
                rgba[bindex * 4 + 3] = 255;
            }
        image_data.push_back(std::move(rgba));
        rtw_image im{};
        im.width = w;
        im.height = h;
        im.rgba = nullptr;  // fixed up in finish()
        images.push_back(im);
        rtw_texture t{};
        t.type = RTW_TEX_IMAGE;
        t.image = (int)images.size() - 1;
        t.odd = t.even = t.perlin = -1;
        texs.push_back(t);
        return (int)texs.size() - 1;
    }

    // ---- materials (material/io*.h assignTo) ----
    int mat(int type, int tex, float param) {
        rtw_material m{};
        m.type = type;
        m.texture = tex;
        m.param = param;
        mats.push_back(m);
        return (int)mats.size() - 1;
    }
    int lambertian(int tex) { return mat(RTW_MAT_LAMBERTIAN, tex, 0.f); }
    int metal(int tex, float fuzz) { return mat(RTW_MAT_METAL, tex, fuzz < 1.f ? fuzz : 1.f); }  // ioMetalMaterial.h:28-32
    int dielectric(float eta) { return mat(RTW_MAT_DIELECTRIC, -1, eta); }
    int diffuse_light(int tex) { return mat(RTW_MAT_DIFFUSE_LIGHT, tex, 0.f); }
    int isotropic(int tex) { return mat(RTW_MAT_ISOTROPIC, tex, 0.f); }

    // ---- primitives ----
    int add(rtw_prim p) { prims.push_back(p); return (int)prims.size() - 1; }
    int sphere(float x, float y, float z, float r, int material, int xf = -1) {
        rtw_prim p{};
        p.type = RTW_PRIM_SPHERE;
        p.material = material;
        p.transform = xf;
        p.v[0] = x; p.v[1] = y; p.v[2] = z; p.v[3] = r;
        return add(p);
    }
    int moving_sphere(float x, float y, float z, float x1, float y1, float z1, float r, float t0, float t1, int material) {
        rtw_prim p{};
        p.type = RTW_PRIM_MOVING_SPHERE;
        p.material = material;
        p.transform = -1;
        p.v[0] = x; p.v[1] = y; p.v[2] = z; p.v[3] = r;
        p.v[4] = x1; p.v[5] = y1; p.v[6] = z1; p.v[7] = t0; p.v[8] = t1;
        return add(p);
    }
    // ioAARect(a0, a1, b0, b1, k, flip, axis)  geometry/ioAARect.h:17-27
    int rect(float a0, float a1, float b0, float b1, float k, bool flip, int axis, int material, int xf = -1) {
        rtw_prim p{};
        p.type = RTW_PRIM_RECT_X + axis;
        p.material = material;
        p.transform = xf;
        p.flip = flip ? 1 : 0;
        p.v[0] = a0; p.v[1] = a1; p.v[2] = b0; p.v[3] = b1; p.v[4] = k;
        return add(p);
    }
    // geometry/ioGeometryGroup.h:27-40
    void box(F3 p0, F3 p1, int material, int xf = -1) {
        rect(p0.x, p1.x, p0.y, p1.y, p0.z, true, 2, material, xf);
        rect(p0.x, p1.x, p0.y, p1.y, p1.z, false, 2, material, xf);
        rect(p0.x, p1.x, p0.z, p1.z, p0.y, true, 1, material, xf);
        rect(p0.x, p1.x, p0.z, p1.z, p1.y, false, 1, material, xf);
        rect(p0.y, p1.y, p0.z, p1.z, p0.x, true, 0, material, xf);
        rect(p0.y, p1.y, p0.z, p1.z, p1.x, false, 0, material, xf);
    }
    int volume_box(F3 p0, F3 p1, float density, int material, int xf) {
        rtw_prim p{};
        p.type = RTW_PRIM_VOLUME_BOX;
        p.material = material;
        p.transform = xf;
        p.v[0] = p0.x; p.v[1] = p0.y; p.v[2] = p0.z; p.v[3] = p1.x; p.v[4] = p1.y; p.v[5] = p1.z; p.v[6] = density;
        return add(p);
    }
    int volume_sphere(F3 c, float r, float density, int material, int xf) {
        rtw_prim p{};
        p.type = RTW_PRIM_VOLUME_SPHERE;
        p.material = material;
        p.transform = xf;
        p.v[0] = c.x; p.v[1] = c.y; p.v[2] = c.z; p.v[3] = r; p.v[4] = density;
        return add(p);
    }
    int xform(const M4& m) { xforms.push_back(make_transform(m)); return (int)xforms.size() - 1; }

    // LightDefinition as the scene functions fill it (e.g. scene/ioScene.h:605-612)
    void light(F3 emission, F3 vecU, F3 vecV, F3 position) {
        rtw_light l{};
        put(l.emission, emission);
        put(l.vecU, vecU);
        put(l.vecV, vecV);
        put(l.position, position);
        l.area = length(cross(vecU, vecV));
        put(l.normal, normalize(cross(vecU, vecV)));
        lights.push_back(l);
    }
    void light_rect(int axis, float a0, float a1, float b0, float b1, float k) {
        desc.has_light_rect = 1;
        desc.light_rect.axis = axis;
        desc.light_rect.a0 = a0; desc.light_rect.a1 = a1;
        desc.light_rect.b0 = b0; desc.light_rect.b1 = b1;
        desc.light_rect.k = k;
    }

    // scene/ioCamera.h:10-34,64-90; lens radius stays 0 (Q3: SysParamter.cameraLensRadius never set);
    // times are Director.cpp:521-522 (cameraTime0 = 0, cameraTime1 = 1 for every scene).
    void camera(F3 from, F3 to, F3 up, float vFov, float aspect, float focus_dist) {
        F3 w = normalize(f3(from.x - to.x, from.y - to.y, from.z - to.z));
        F3 u = normalize(cross(up, w));
        F3 v = cross(w, u);
        float theta = vFov * kPi / 180.0f;
        float halfHeight = tanf(theta / 2.0f);
        float halfWidth = aspect * halfHeight;
        F3 ll = from - halfWidth * focus_dist * u - halfHeight * focus_dist * v - focus_dist * w;
        F3 horizontal = 2.0f * halfWidth * focus_dist * u;
        F3 vertical = 2.0f * halfHeight * focus_dist * v;
        rtw_camera& c = desc.camera;
        put(c.origin, from);
        put(c.u, u);
        put(c.v, v);
        put(c.w, w);
        put(c.lower_left, ll);
        put(c.horizontal, horizontal);
        put(c.vertical, vertical);
        c.lens_radius = 0.f;
        c.time0 = 0.f;
        c.time1 = 1.f;
    }

    void finish() {
        for (size_t i = 0; i < images.size(); i++) images[i].rgba = image_data[i].data();
        desc.n_prims = (int)prims.size();       desc.prims = prims.data();
        desc.n_transforms = (int)xforms.size(); desc.transforms = xforms.empty() ? nullptr : xforms.data();
        desc.n_materials = (int)mats.size();    desc.materials = mats.data();
        desc.n_textures = (int)texs.size();     desc.textures = texs.empty() ? nullptr : texs.data();
        desc.n_perlin = (int)perlin.size();     desc.perlin = perlin.empty() ? nullptr : perlin.data();
        desc.n_images = (int)images.size();     desc.images = images.empty() ? nullptr : images.data();
        desc.n_lights = (int)lights.size();     desc.lights = lights.empty() ? nullptr : lights.data();
        desc.sky_light = lights.empty() ? 1 : 0;  // Director.cpp:523
    }

    // The IOW small-sphere field shared by scenes 1 and 2 (scene/ioScene.h:200-252, :373-417).
    void small_spheres(uint32_t seed, bool moving) {
        for (int a = -11; a < 11; a++) {
            for (int b = -11; b < 11; b++) {
                float chooseMat = randf(seed);
                float x = a + 0.8f * randf(seed);
                float y = 0.2f;
                float z = b + 0.9f * randf(seed);
                float z_squared = (z) * (z);
                float dist = sqrtf((x - 4.0f) * (x - 4.0f) + z_squared);
                if ((dist > 0.9f) || ((z_squared > 0.7f) && ((x * x - 16.0f) > -2.f))) {
                    if (chooseMat < 0.70f) {
                        // make_float3(randf, randf, randf): right-to-left (Q33)
                        float cz = randf(seed), cy = randf(seed), cx = randf(seed);
                        int m = lambertian(tex_constant(f3(cx, cy, cz)));
                        if (moving) moving_sphere(x, y, z, x, y + 0.18f, z, 0.2f, 0.f, 1.f, m);
                        else sphere(x, y, z, 0.2f, m);
                    } else if (chooseMat < 0.85f) {
                        // ioMetalMaterial(ioConstantTexture(make_float3(...)), 0.5f*randf): right-to-left
                        float fuzz = 0.5f * randf(seed);
                        float cz = 0.5f * (1.0f - randf(seed));
                        float cy = 0.5f * (1.0f - randf(seed));
                        float cx = 0.5f * (1.0f - randf(seed));
                        int m = metal(tex_constant(f3(cx, cy, cz)), fuzz);
                        sphere(x, y, z, 0.2f, m);
                    } else if (chooseMat < 0.93f) {
                        sphere(x, y, z, 0.2f, dielectric(1.5f));
                    } else {
                        sphere(x, y, z, 0.2f, dielectric(1.5f));
                        sphere(x, y, z, (0.2f - 0.007f), dielectric(1.5f));
                    }
                }
            }
        }
    }

    // ---- scene 0: scene/ioScene.h:495-629 ----
    void cornell_box(int Nx, int Ny) {
        description = "Cornell box";
        int wallRed = lambertian(tex_constant(f3(0.65f, 0.05f, 0.05f)));
        int wallGreen = lambertian(tex_constant(f3(0.12f, 0.45f, 0.15f)));
        int wallWhite = lambertian(tex_constant(f3(0.73f, 0.73f, 0.73f)));
        int aluminum = metal(tex_constant(f3((float)0.91, (float)0.92, (float)0.92)), (float)0.018);
        int light15 = tex_constant(f3(15.f, 15.f, 15.f));
        sphere(190.f, 90.f, 190.f, 90.f, dielectric(1.5f));
        rect(0.f, 555.f, 0.f, 555.f, 555.f, true, 0, wallGreen);
        rect(0.f, 555.f, 0.f, 555.f, 0.f, false, 0, wallRed);
        rect(0.f, 555.f, 0.f, 555.f, 555.f, true, 1, wallWhite);
        rect(0.f, 555.f, 0.f, 555.f, 0.f, false, 1, wallWhite);
        rect(0.f, 555.f, 0.f, 555.f, 555.f, true, 2, wallWhite);
        rect(213.f, 343.f, 227.f, 332.f, 554.9f, true, 1, diffuse_light(light15));
        F3 b1size = f3(165.f, 330.f, 165.f);
        F3 b1tr = f3(265.f, 0.f, 295.f);
        M4 transf = mul(translate(b1tr), rotateY(15.f));
        box(f3(0.f, 0.f, 0.f), b1size, aluminum, xform(transf));
        light(f3(15.f, 15.f, 15.f), f3(343.f - 213.f, 0, 0.f), f3(0, 0, 332.f - 227.f), f3(213.f, 554.f, 227.f));
        light_rect(1, 213.f, 343.f, 227.f, 332.f, 554.9f);  // ioScene.h:112
        camera(f3(278.f, 278.f, -800.f), f3(278.f, 278.f, 0.f), f3(0.0f, 1.0f, 0.0f), 40.0f, float(Nx) / float(Ny), 10.f);
    }

    // ---- scene 1: scene/ioScene.h:180-304 ----
    void moving_spheres(int Nx, int Ny) {
        description = "InOneWeekend final scene with moving spheres";
        int fiftyPercentGrey = tex_constant(f3(0.5f, 0.5f, 0.5f));
        int fiftyPercentReddishGrey = tex_constant(f3(0.7f, 0.6f, 0.5f));
        int reddish = tex_constant(f3(0.4f, 0.2f, 0.1f));
        sphere(0.0f, -1000.0f, 0.0, 1000.0f, lambertian(fiftyPercentGrey));
        int m1 = dielectric(1.5f), m2 = lambertian(reddish), m3 = metal(fiftyPercentReddishGrey, 0.1f);
        sphere(0.0f, 1.0f, 0.0, 1.0f, m1);
        sphere(-4.0f, 1.0f, 0.0, 1.0f, m2);
        sphere(4.0f, 1.0f, 0.0, 1.0f, m3);
        small_spheres(0x314759, true);
        camera(f3(13.0f, 2.0f, 3.0f), f3(0.0f, 0.0f, 0.0f), f3(0.0f, 1.0f, 0.0f), 20.0f, float(Nx) / float(Ny), 10.f);
    }

    // ---- scene 2: scene/ioScene.h:308-493 ----
    bool in_one_weekend_light(int Nx, int Ny) {
        description = "IOW Scene with a light box";
        int constantGrey = tex_constant(f3(0.7f, 0.7f, 0.7f));
        int light16 = tex_constant(f3(16.f, 16.f, 16.f));
        // material list order decides the Perlin draw order: the noise material is first (ioScene.h:338)
        int noise1 = tex_noise(1.f);
        sphere(0.0f, -1000.0f, 0.0, 1000.0f, lambertian(noise1));
        int earth = tex_image("earthmap.ppm");
        if (earth < 0) return false;
        int mA = metal(constantGrey, 0.4f), mB = lambertian(earth), mC = dielectric(1.5f);
        sphere(-4.0f, 1.0f, 0.0, 1.0f, mA);
        sphere(0.0f, 1.0f, 0.0, 1.0f, mB);
        sphere(4.0f, 1.0f, 0.0, 1.0f, mC);
        rect(3.f, 5.f, 1.f, 3.f, -2.0f, false, 2, diffuse_light(light16));
        light(f3(16.f, 16.f, 16.f), f3(5.f - 3.f, 0, 0.f), f3(0, 3.f - 1.f, 0.f), f3(3.f, 1.f, -2.f));
        small_spheres(0x6314759, false);
        light_rect(2, 3.f, 5.f, 2.3f, 3.f + 3.f, -2.0f);  // ioScene.h:124
        camera(f3(13.0f, 2.0f, 3.0f), f3(0.0f, 0.0f, 0.0f), f3(0.0f, 1.0f, 0.0f), 20.0f, float(Nx) / float(Ny), 10.f);
        return true;
    }

    // ---- scene 3: scene/ioScene.h:632-788 (no LightDefinition is pushed → sky light, no NEE) ----
    void volumes_cornell_box(int Nx, int Ny) {
        description = "Cornell box with volumes (participating media)";
        int wallRed = lambertian(tex_constant(f3(0.65f, 0.05f, 0.05f)));
        int wallGreen = lambertian(tex_constant(f3(0.12f, 0.45f, 0.15f)));
        int wallWhite = lambertian(tex_constant(f3(0.73f, 0.73f, 0.73f)));
        int light15 = tex_constant(f3(15.f, 15.f, 15.f));
        int blackFog = isotropic(tex_constant(f3(0.f, 0.f, 0.f)));
        int whiteFog = isotropic(tex_constant(f3(1.f, 1.f, 1.f)));
        rect(0.f, 555.f, 0.f, 555.f, 555.f, true, 0, wallGreen);
        rect(0.f, 555.f, 0.f, 555.f, 0.f, false, 0, wallRed);
        rect(0.f, 555.f, 0.f, 555.f, 555.f, true, 1, wallWhite);
        rect(0.f, 555.f, 0.f, 555.f, 0.f, false, 1, wallWhite);
        rect(0.f, 555.f, 0.f, 555.f, 555.f, true, 2, wallWhite);
        rect(213.f, 343.f, 227.f, 332.f, 554.f, true, 1, diffuse_light(light15));
        const float z1Theta = -12.5f * (kPi / 180.f);
        F3 b1size = f3(165.f, 330.f, 165.f);
        F3 b1tr = f3(265.f, fabsf(sinf(z1Theta)) * b1size.x + 0.f, 255.f);
        M4 trans = mul(mul(translate(b1tr), rotateZ(z1Theta * (180.f / kPi))), rotateY(15.f));
        volume_box(f3(0.f, 0.f, 0.f), b1size, 0.006f, blackFog, xform(trans));
        F3 b2origin = f3(0.f, 0.f, 0.f) + f3(165.f / 2.f, 75.f, 165.f / 2.f);
        F3 b2tr = f3(130.f, 0.f, 65.f);
        volume_sphere(b2origin, 75.f, 0.005f, whiteFog, xform(translate(b2tr)));
        light_rect(1, 213.f, 343.f, 227.f, 332.f, 554.f);  // ioScene.h:133 (unused: numLights == 0)
        camera(f3(278.f, 278.f, -800.f), f3(278.f, 278.f, 0.f), f3(0.0f, 1.0f, 0.0f), 40.0f, float(Nx) / float(Ny), 10.f);
    }

    // ---- scene 4: scene/ioScene.h:791-982 (prim order == instanceId order) ----
    bool the_next_week_final(int Nx, int Ny) {
        description = "The Next Week final scene";
        int brown = tex_constant(f3(0.7f, 0.3f, 0.1f));
        int groundGreenish = tex_constant(f3(0.48f, 0.83f, 0.53f));
        int metal1 = tex_constant(f3(0.8f, 0.8f, 0.9f));
        int light7 = tex_constant(f3(7.f, 7.f, 7.f));
        uint32_t seed = 0x6314759;
        // id 0: light
        rect(123.f, 423.f, 147.f, 412.f, 554.f, true, 1, diffuse_light(light7));
        light(f3(7.f, 7.f, 7.f), f3(423.f - 123.f, 0, 0.f), f3(0, 0.f, 412.f - 147.f), f3(123.f, 554.f, 147.f));
        F3 center = f3(400.f, 400.f, 200.f);
        F3 center1tr = center + f3(30.f, 0.f, 0.f);
        sphere(260.f, 150.f, 45.f, 50.f, dielectric(1.5f));                        // id 1
        sphere(0.f, 150.f, 145.f, 50.f, metal(metal1, 0.2f));                      // id 2
        F3 centerGlassy = f3(360.f, 150.f, 45.f);
        sphere(centerGlassy.x, centerGlassy.y, centerGlassy.z, 70.f, dielectric(1.5f));  // id 3
        sphere(0.f, 0.f, 0.f, 5000.f, dielectric(1.5f));                           // id 4 (Q34)
        int earth = tex_image("earthmap.ppm");
        if (earth < 0) return false;
        int mEarth = lambertian(earth);
        // Perlin tables are drawn when the materials are flattened, in material order:
        // only noisep1 (material 6) is referenced (ioScene.h:842-843).
        int noisep1 = tex_noise(0.1f);
        int mNoise = lambertian(noisep1);
        sphere(400.f, 200.f, 400.f, 100.0f, mEarth);                               // id 5
        sphere(220.f, 280.f, 300.f, 80.f, mNoise);                                 // id 6
        moving_sphere(center.x, center.y, center.z, center1tr.x, center1tr.y, center1tr.z, 50.f, 0.f, 1.f,
                      lambertian(brown));                                          // id 7
        int ground = lambertian(groundGreenish);
        for (int i = 0; i < 20; i++) {
            for (int j = 0; j < 20; j++) {
                float w = 100.f;
                float x0 = -1000 + i * w;
                float z0 = -1000 + j * w;
                float y0 = 0.f;
                float x1 = x0 + w;
                float y1 = 100 * (randf(seed) + 0.01f);
                float z1 = z0 + w;
                box(f3(x0, y0, z0), f3(x1, y1, z1), ground);                       // ids 8 + (i*20+j)*6 + jj
            }
        }
        int glassyBlueFog = isotropic(tex_constant(f3(0.2f, 0.4f, 0.9f)));
        int ambientFog = isotropic(tex_constant(f3(0.95f, 0.95f, 0.95f)));
        volume_sphere(centerGlassy, 70.f, 0.2f, glassyBlueFog, -1);                // id 2408
        volume_sphere(f3(0.f, 0.f, 0.f), 500.f, 8e-5f, ambientFog, -1);            // id 2409
        int white = lambertian(tex_constant(f3(0.93f, 0.93f, 0.93f)));
        M4 transmat = mul(translate(f3(-100.f, 270.f, 395.f)), rotateY(20.f));
        int xf = xform(transmat);
        for (int j = 0; j < 1000; j++) {
            // make_float3(165*randf, 165*randf, 165*randf): right-to-left (Q33)
            float cz = 165 * randf(seed), cy = 165 * randf(seed), cx = 165 * randf(seed);
            sphere(cx, cy, cz, 10.f, white, xf);                                   // ids 2410..3409 (Q21)
        }
        light_rect(1, 123.f, 423.f, 147.f, 412.f, 554.f);  // ioScene.h:142
        camera(f3(478.f, 278.f, -600.f), f3(278.f, 278.f, 0.f), f3(0.f, 1.f, 0.f), 40.0f, (float(Nx) / float(Ny)), 10.f);
        return true;
    }

    // ---- scene 5: synthetic stress scene C5 (no reference counterpart) ----
    void stress(int Nx, int Ny, int n) {
        description = "Synthetic stress scene (C5)";
        uint32_t seed = 0x6314759;
        std::vector<int> xf(1024);
        for (int k = 0; k < 1024; k++) xf[k] = xform(mul(translate(f3(0.f, (float)k, 0.f)), rotateY((float)k * 0.35f)));
        int metalTex = tex_constant(f3(0.8f, 0.8f, 0.8f));
        int mMetal = metal(metalTex, 0.1f);
        int mGlass = dielectric(1.5f);
        for (int i = 0; i < n; i++) {
            float cx = 1000.f * randf(seed), cy = 1000.f * randf(seed), cz = 1000.f * randf(seed);
            float sz = 0.5f + 1.5f * randf(seed);
            int mk = i % 20;
            int m;
            if (mk < 14) {
                float r = randf(seed), g = randf(seed), b = randf(seed);
                m = lambertian(tex_constant(f3(r, g, b)));
            } else if (mk < 17) m = mMetal;
            else m = mGlass;
            int t = xf[i % 1024];
            if ((i & 1) == 0) sphere(cx, cy, cz, sz, m, t);
            else {
                int axis = i % 3;
                float c[3] = {cx, cy, cz};
                float k = c[axis];
                float a = c[axis == 0 ? 1 : 0], b = c[axis == 2 ? 1 : 2];
                rect(a - sz, a + sz, b - sz, b + sz, k, false, axis, m, t);
            }
        }
        camera(f3(1500.f, 800.f, -1500.f), f3(500.f, 500.f, 500.f), f3(0.f, 1.f, 0.f), 40.0f, float(Nx) / float(Ny), 10.f);
    }
};

extern "C" int rtw_scene_create(int scene_id, int Nx, int Ny, const char* asset_dir, int n_stress_prims,
                                rtw_scene** out) {
    if (!out || Nx <= 0 || Ny <= 0) return RTW_ERR_INVALID;
    *out = nullptr;
    rtw_scene* s = new rtw_scene();
    s->asset_dir = asset_dir ? asset_dir : "assets";
    bool ok = true;
    switch (scene_id) {  // scene/ioScene.h:74-93
    case 0: s->cornell_box(Nx, Ny); break;
    case 1: s->moving_spheres(Nx, Ny); break;
    case 2: ok = s->in_one_weekend_light(Nx, Ny); break;
    case 3: s->volumes_cornell_box(Nx, Ny); break;
    case 4: ok = s->the_next_week_final(Nx, Ny); break;
    case 5: s->stress(Nx, Ny, n_stress_prims > 0 ? n_stress_prims : (1 << 20)); break;
    default:
        fprintf(stderr, "ERROR: Scene %d unknown.\n", scene_id);  // ioScene.h:95
        delete s;
        return RTW_ERR_INVALID;
    }
    if (!ok) {
        fprintf(stderr, "ERROR: %s\n", s->error.c_str());
        delete s;
        return RTW_ERR_IO;
    }
    s->finish();
    *out = s;
    return RTW_OK;
}

extern "C" const rtw_scene_desc* rtw_scene_get_desc(const rtw_scene* s) { return s ? &s->desc : nullptr; }
extern "C" const char* rtw_scene_description(const rtw_scene* s) { return s ? s->description.c_str() : ""; }
extern "C" void rtw_scene_free(rtw_scene* s) { delete s; }

// ---- scene wire format (the step BEFORE the path) ---------------------------
// The reference has no scene file: its five scenes are C++ functions (scene/ioScene.h:180-982) that
// fill OptiX records.  Here a flattened scene (rtw_scene_desc) is data: one little-endian file,
//   u32 magic "RTWS", u32 version, u32 header bytes, then the header
//   { i32 n_prims, n_transforms, n_materials, n_textures, n_perlin, n_images, n_lights, has_light_rect,
//     sky_light, reserved; rtw_light_rect; rtw_camera; u32 description bytes },
//   the description text, the arrays in that order as their C-ABI PODs (include/rtw.h) — every image as
//   { i32 width, height } + RGBA8 texels —, and a u64 FNV-1a checksum of everything before it.
// rtw_scene_load validates counts, sizes, the checksum and every index a kernel would follow, so a file
// that loads cannot send the device out of bounds.
namespace {

constexpr uint32_t kSceneMagic = 0x53575452u;   // "RTWS"
constexpr uint32_t kSceneVersion = 1u;

struct SceneFileHeader {
    int32_t n_prims, n_transforms, n_materials, n_textures, n_perlin, n_images, n_lights, has_light_rect;
    int32_t sky_light, reserved;
    rtw_light_rect light_rect;
    rtw_camera camera;
    uint32_t description_bytes;
};

struct Fnv {
    uint64_t h = 1469598103934665603ull;
    void add(const void* p, size_t n) {
        const uint8_t* b = static_cast<const uint8_t*>(p);
        for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ull; }
    }
};

struct SceneWriter {
    FILE* f;
    Fnv sum;
    bool ok = true;
    void put(const void* p, size_t n) {
        if (n == 0) return;
        if (fwrite(p, 1, n, f) != n) ok = false;
        sum.add(p, n);
    }
};

struct SceneReader {
    FILE* f;
    Fnv sum;
    bool ok = true;
    void get(void* p, size_t n) {
        if (n == 0) return;
        if (!ok || fread(p, 1, n, f) != n) { ok = false; memset(p, 0, n); return; }
        sum.add(p, n);
    }
};

bool index_ok(int32_t i, int32_t n, bool may_be_none) { return (may_be_none && i == -1) || (i >= 0 && i < n); }

// every index and enumeration a kernel follows
const char* validate_desc(const rtw_scene_desc& d) {
    if (d.n_prims < 1 || d.n_prims > (1 << 25)) return "primitive count out of range [1, 2^25]";
    if (d.n_materials < 1) return "no materials";
    if (d.n_lights < 0 || d.n_lights > 8) return "light count out of range [0, 8]";
    if (d.n_transforms < 0 || d.n_textures < 0 || d.n_perlin < 0 || d.n_images < 0) return "negative table size";
    if (!d.prims || !d.materials || (d.n_transforms && !d.transforms) || (d.n_textures && !d.textures) || (d.n_perlin && !d.perlin) ||
        (d.n_images && !d.images) || (d.n_lights && !d.lights))
        return "table pointer missing";
    for (int i = 0; i < d.n_prims; i++) {
        const rtw_prim& p = d.prims[i];
        if (p.type < 0 || p.type >= RTW_PRIM_TYPE_COUNT) return "primitive type out of range";
        if (!index_ok(p.material, d.n_materials, false)) return "primitive material index out of range";
        if (!index_ok(p.transform, d.n_transforms, true)) return "primitive transform index out of range";
        // A moving sphere's LBVH box covers its centres for gather times inside [time0, time1] (device/lbvh.h);
        // moving_center() extrapolates outside, so the camera's shutter interval must lie inside the sphere's.
        if (p.type == RTW_PRIM_MOVING_SPHERE && p.v[7] != p.v[8]) {
            const float lo = p.v[7] < p.v[8] ? p.v[7] : p.v[8], hi = p.v[7] < p.v[8] ? p.v[8] : p.v[7];
            const float c0 = d.camera.time0 < d.camera.time1 ? d.camera.time0 : d.camera.time1;
            const float c1 = d.camera.time0 < d.camera.time1 ? d.camera.time1 : d.camera.time0;
            if (!(c0 >= lo && c1 <= hi)) return "camera shutter interval outside a moving sphere's time range";
        }
    }
    for (int i = 0; i < d.n_materials; i++) {
        const rtw_material& m = d.materials[i];
        if (m.type < 0 || m.type >= RTW_MAT_TYPE_COUNT) return "material type out of range";
        if (!index_ok(m.texture, d.n_textures, true)) return "material texture index out of range";
        const bool textured = m.type == RTW_MAT_LAMBERTIAN || m.type == RTW_MAT_DIFFUSE_LIGHT || m.type == RTW_MAT_METAL ||
                              m.type == RTW_MAT_ISOTROPIC;
        if (textured && m.texture < 0) return "textured material without a texture";
    }
    for (int i = 0; i < d.n_textures; i++) {
        const rtw_texture& t = d.textures[i];
        if (t.type < 0 || t.type >= RTW_TEX_TYPE_COUNT) return "texture type out of range";
        if (t.type == RTW_TEX_CHECKER && (!index_ok(t.odd, d.n_textures, false) || !index_ok(t.even, d.n_textures, false)))
            return "checker child index out of range";
        if (t.type == RTW_TEX_CHECKER && (d.textures[t.odd].type == RTW_TEX_CHECKER || d.textures[t.even].type == RTW_TEX_CHECKER))
            return "nested checker textures";
        if (t.type == RTW_TEX_NOISE && !index_ok(t.perlin, d.n_perlin, false)) return "noise table index out of range";
        if (t.type == RTW_TEX_IMAGE && !index_ok(t.image, d.n_images, false)) return "image index out of range";
    }
    for (int i = 0; i < d.n_perlin; i++)
        for (int k = 0; k < 256; k++) {
            const rtw_perlin& p = d.perlin[i];
            if ((p.perm_x[k] | p.perm_y[k] | p.perm_z[k]) & ~255) return "Perlin permutation entry out of range";
        }
    for (int i = 0; i < d.n_images; i++)
        if (d.images[i].width < 1 || d.images[i].height < 1 || !d.images[i].rgba) return "empty image";
    if (d.has_light_rect && (d.light_rect.axis < 0 || d.light_rect.axis > 2)) return "light rectangle axis out of range";
    return nullptr;
}

}  // namespace

// the one strict validator, also behind rtw_upload_scene (api/rtw_api.cu)
const char* rtw_internal_validate_desc(const rtw_scene_desc* d) { return d ? validate_desc(*d) : "null scene description"; }

static std::string g_scene_io_error;
extern "C" const char* rtw_scene_io_error(void) { return g_scene_io_error.c_str(); }

extern "C" int rtw_scene_save(const rtw_scene_desc* d, const char* description, const char* path) {
    if (!d || !path) return RTW_ERR_INVALID;
    if (const char* why = validate_desc(*d)) {
        g_scene_io_error = why;
        return RTW_ERR_INVALID;
    }
    FILE* f = fopen(path, "wb");
    if (!f) {
        g_scene_io_error = std::string("cannot open ") + path;
        return RTW_ERR_IO;
    }
    SceneWriter w{f};
    const std::string text = description ? description : "";
    SceneFileHeader h{};
    h.n_prims = d->n_prims; h.n_transforms = d->n_transforms; h.n_materials = d->n_materials;
    h.n_textures = d->n_textures; h.n_perlin = d->n_perlin; h.n_images = d->n_images; h.n_lights = d->n_lights;
    h.has_light_rect = d->has_light_rect; h.sky_light = d->sky_light; h.reserved = 0;
    h.light_rect = d->light_rect;
    h.camera = d->camera;
    h.description_bytes = (uint32_t)text.size();
    const uint32_t pre[3] = {kSceneMagic, kSceneVersion, (uint32_t)sizeof(SceneFileHeader)};
    w.put(pre, sizeof(pre));
    w.put(&h, sizeof(h));
    w.put(text.data(), text.size());
    w.put(d->prims, sizeof(rtw_prim) * (size_t)d->n_prims);
    w.put(d->transforms, sizeof(rtw_transform) * (size_t)d->n_transforms);
    w.put(d->materials, sizeof(rtw_material) * (size_t)d->n_materials);
    w.put(d->textures, sizeof(rtw_texture) * (size_t)d->n_textures);
    w.put(d->perlin, sizeof(rtw_perlin) * (size_t)d->n_perlin);
    for (int i = 0; i < d->n_images; i++) {
        const int32_t wh[2] = {d->images[i].width, d->images[i].height};
        w.put(wh, sizeof(wh));
        w.put(d->images[i].rgba, (size_t)wh[0] * (size_t)wh[1] * 4);
    }
    w.put(d->lights, sizeof(rtw_light) * (size_t)d->n_lights);
    const uint64_t sum = w.sum.h;
    if (fwrite(&sum, 1, sizeof(sum), f) != sizeof(sum)) w.ok = false;
    if (fclose(f) != 0) w.ok = false;
    if (!w.ok) {
        g_scene_io_error = std::string("short write to ") + path;
        return RTW_ERR_IO;
    }
    return RTW_OK;
}

extern "C" int rtw_scene_load(const char* path, rtw_scene** out) {
    if (!path || !out) return RTW_ERR_INVALID;
    *out = nullptr;
    FILE* f = fopen(path, "rb");
    if (!f) {
        g_scene_io_error = std::string("cannot open ") + path;
        return RTW_ERR_IO;
    }
    fseek(f, 0, SEEK_END);
    const long file_bytes = ftell(f);
    fseek(f, 0, SEEK_SET);
    SceneReader r{f};
    rtw_scene* s = new rtw_scene();
    auto bad = [&](int code, const std::string& why) {
        g_scene_io_error = why;
        fclose(f);
        delete s;
        return code;
    };
    uint32_t pre[3];
    r.get(pre, sizeof(pre));
    if (!r.ok || pre[0] != kSceneMagic) return bad(RTW_ERR_INVALID, "not an RTWS scene file");
    if (pre[1] != kSceneVersion || pre[2] != sizeof(SceneFileHeader)) return bad(RTW_ERR_INVALID, "unsupported scene file version");
    SceneFileHeader h;
    r.get(&h, sizeof(h));
    if (!r.ok) return bad(RTW_ERR_IO, "truncated scene file (header)");
    const int32_t counts[7] = {h.n_prims, h.n_transforms, h.n_materials, h.n_textures, h.n_perlin, h.n_images, h.n_lights};
    for (int32_t n : counts)
        if (n < 0 || n > (1 << 25)) return bad(RTW_ERR_INVALID, "array count out of range");
    // the arrays must fit in what is left of the file (a count cannot make us allocate more than the file holds)
    const unsigned long long need = (unsigned long long)h.description_bytes + sizeof(rtw_prim) * (unsigned long long)h.n_prims +
        sizeof(rtw_transform) * (unsigned long long)h.n_transforms + sizeof(rtw_material) * (unsigned long long)h.n_materials +
        sizeof(rtw_texture) * (unsigned long long)h.n_textures + sizeof(rtw_perlin) * (unsigned long long)h.n_perlin +
        8ull * (unsigned long long)h.n_images + sizeof(rtw_light) * (unsigned long long)h.n_lights + 8ull;
    if (file_bytes < 0 || need + sizeof(pre) + sizeof(h) > (unsigned long long)file_bytes) return bad(RTW_ERR_IO, "truncated scene file");
    s->description.resize(h.description_bytes);
    r.get(&s->description[0], h.description_bytes);
    s->prims.resize(h.n_prims);       r.get(s->prims.data(), sizeof(rtw_prim) * s->prims.size());
    s->xforms.resize(h.n_transforms); r.get(s->xforms.data(), sizeof(rtw_transform) * s->xforms.size());
    s->mats.resize(h.n_materials);    r.get(s->mats.data(), sizeof(rtw_material) * s->mats.size());
    s->texs.resize(h.n_textures);     r.get(s->texs.data(), sizeof(rtw_texture) * s->texs.size());
    s->perlin.resize(h.n_perlin);     r.get(s->perlin.data(), sizeof(rtw_perlin) * s->perlin.size());
    for (int i = 0; i < h.n_images && r.ok; i++) {
        int32_t wh[2] = {0, 0};
        r.get(wh, sizeof(wh));
        if (!r.ok || wh[0] < 1 || wh[1] < 1 || (unsigned long long)wh[0] * (unsigned long long)wh[1] * 4ull > (unsigned long long)file_bytes)
            return bad(RTW_ERR_INVALID, "bad image size");
        rtw_image im{};
        im.width = wh[0];
        im.height = wh[1];
        s->images.push_back(im);
        s->image_data.emplace_back((size_t)wh[0] * (size_t)wh[1] * 4);
        r.get(s->image_data.back().data(), s->image_data.back().size());
    }
    s->lights.resize(h.n_lights);     r.get(s->lights.data(), sizeof(rtw_light) * s->lights.size());
    if (!r.ok) return bad(RTW_ERR_IO, "truncated scene file");
    uint64_t sum = 0;
    if (fread(&sum, 1, sizeof(sum), f) != sizeof(sum)) return bad(RTW_ERR_IO, "truncated scene file (checksum)");
    if (sum != r.sum.h) return bad(RTW_ERR_INVALID, "scene file checksum mismatch");
    s->desc.has_light_rect = h.has_light_rect;
    s->desc.light_rect = h.light_rect;
    s->desc.camera = h.camera;
    s->finish();
    s->desc.sky_light = h.sky_light;
    if (const char* why = validate_desc(s->desc)) return bad(RTW_ERR_INVALID, why);
    fclose(f);
    *out = s;
    return RTW_OK;
}

// Director::printPPM, Director.cpp:1010-1031 (source buffer is the un-denoised frame, Q30)
extern "C" int rtw_write_ppm(const char* path, int Nx, int Ny, const float* rgba) {
    if (!rgba || Nx <= 0 || Ny <= 0) return RTW_ERR_INVALID;
    FILE* f = (!path || strcmp(path, "-") == 0) ? stdout : fopen(path, "w");
    if (!f) return RTW_ERR_IO;
    fprintf(f, "P3\n%d %d\n255\n", Nx, Ny);
    // sutil clamp(f, a, b) = fmaxf(a, fminf(f, b))
    auto clamp01 = [](float v) { return fmaxf(0.f, fminf(v, 1.f)); };
    for (int j = Ny - 1; j >= 0; j--)
        for (int i = 0; i < Nx; i++) {
            const float* p = rgba + ((size_t)Nx * j + i) * 4;
            int ir = int(255.99f * clamp01(p[0]));
            int ig = int(255.99f * clamp01(p[1]));
            int ib = int(255.99f * clamp01(p[2]));
            fprintf(f, "%d %d %d\n", ir, ig, ib);
        }
    if (f != stdout) fclose(f);
    else fflush(f);
    return RTW_OK;
}
