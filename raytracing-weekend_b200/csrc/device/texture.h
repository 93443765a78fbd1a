// texture.h — texture callables of the reference as one switch.
//   texture/constantTexture.cu:5-10, nullTexture.cu:7-12, checkeredTexture.cu:9-20,
//   imageTexture.cu:10-16, noiseTexture.cu:19-83
#pragma once
#include "detmath.h"
#include "layout.h"

namespace rtw {

RTW_HD float4 ldg4(const float4* p) {
#if defined(__CUDA_ARCH__)
    return __ldg(p);
#else
    return *p;
#endif
}
RTW_HD int ldgi(const int* p) {
#if defined(__CUDA_ARCH__)
    return __ldg(p);
#else
    return *p;
#endif
}

// texture/noiseTexture.cu:19-37
RTW_HD float perlin_interp(const float3 c[2][2][2], float u, float v, float w) {
    float uu = u * u * (3 - 2 * u);
    float vv = v * v * (3 - 2 * v);
    float ww = w * w * (3 - 2 * w);
    float accum = 0;
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
        for (int j = 0; j < 2; j++)
#pragma unroll
            for (int k = 0; k < 2; k++) {
                float3 weight_v = f3(u - i, v - j, w - k);
                accum += (i * uu + (1 - i) * (1 - uu)) * (j * vv + (1 - j) * (1 - vv)) *
                         (k * ww + (1 - k) * (1 - ww)) * dot(c[i][j][k], weight_v);
            }
    return accum;
}

// texture/noiseTexture.cu:39-55
RTW_HD float perlin_noise(const DScene& sc, int table, float3 p) {
    const float4* ranvec = sc.perlin_vec + 256 * table;
    const int* perm_x = sc.perlin_perm + 768 * table;
    const int* perm_y = perm_x + 256;
    const int* perm_z = perm_x + 512;
    float u = p.x - floorf(p.x);
    float v = p.y - floorf(p.y);
    float w = p.z - floorf(p.z);
    int i = (int)floorf(p.x);
    int j = (int)floorf(p.y);
    int k = (int)floorf(p.z);
    float3 c[2][2][2];
#pragma unroll
    for (int di = 0; di < 2; di++)
#pragma unroll
        for (int dj = 0; dj < 2; dj++)
#pragma unroll
            for (int dk = 0; dk < 2; dk++) {
                int idx = ldgi(perm_x + ((i + di) & 255)) ^ ldgi(perm_y + ((j + dj) & 255)) ^ ldgi(perm_z + ((k + dk) & 255));
                c[di][dj][dk] = f3(ldg4(ranvec + idx));
            }
    return perlin_interp(c, u, v, w);
}

// texture/noiseTexture.cu:57-70
RTW_HD float perlin_turb(const DScene& sc, int table, float3 p) {
    float accum = 0;
    float3 temp_p = p;
    float weight = 1.0f;
    for (int i = 0; i < 7; i++) {
        accum += weight * perlin_noise(sc, table, temp_p);
        weight *= 0.5f;
        temp_p = temp_p * 2.f;
    }
    return fabsf(accum);
}

// The reference samples a cudaTextureObject (normalised coords, clamp, linear;
// texture/ioTexture.h:284-300).  Q20 NORMALISED: manual fp32 bilinear on the
// RGBA8 texels, x = u*W - 0.5, NaN → 0, indices clamped (DESIGN.md "image texture").
RTW_HD float3 image_lookup(const DScene& sc, int image, float u, float v) {
    const DImage im = sc.images[image];
    float x = u * (float)im.width - 0.5f;
    float y = v * (float)im.height - 0.5f;
    if (!(x == x)) x = 0.f;
    if (!(y == y)) y = 0.f;
    float fx0 = floorf(x), fy0 = floorf(y);
    float fx = x - fx0, fy = y - fy0;
    float wmax = (float)(im.width - 1), hmax = (float)(im.height - 1);
    int x0 = (int)fminf(fmaxf(fx0, 0.f), wmax), x1 = (int)fminf(fmaxf(fx0 + 1.f, 0.f), wmax);
    int y0 = (int)fminf(fmaxf(fy0, 0.f), hmax), y1 = (int)fminf(fmaxf(fy0 + 1.f, 0.f), hmax);
    uchar4 p00, p10, p01, p11;
#if defined(__CUDA_ARCH__)
    p00 = __ldg(im.texels + (size_t)y0 * im.width + x0);
    p10 = __ldg(im.texels + (size_t)y0 * im.width + x1);
    p01 = __ldg(im.texels + (size_t)y1 * im.width + x0);
    p11 = __ldg(im.texels + (size_t)y1 * im.width + x1);
#else
    p00 = im.texels[(size_t)y0 * im.width + x0];
    p10 = im.texels[(size_t)y0 * im.width + x1];
    p01 = im.texels[(size_t)y1 * im.width + x0];
    p11 = im.texels[(size_t)y1 * im.width + x1];
#endif
    float3 t00 = f3((float)p00.x / 255.0f, (float)p00.y / 255.0f, (float)p00.z / 255.0f);
    float3 t10 = f3((float)p10.x / 255.0f, (float)p10.y / 255.0f, (float)p10.z / 255.0f);
    float3 t01 = f3((float)p01.x / 255.0f, (float)p01.y / 255.0f, (float)p01.z / 255.0f);
    float3 t11 = f3((float)p11.x / 255.0f, (float)p11.y / 255.0f, (float)p11.z / 255.0f);
    float3 r0 = t00 * (1.0f - fx) + t10 * fx;
    float3 r1 = t01 * (1.0f - fx) + t11 * fx;
    return r0 * (1.0f - fy) + r1 * fy;
}

// float3 tex(MaterialParams const&, u, v, p): one switch instead of direct callables.
RTW_HDN inline float3 texture_eval(const DScene& sc, int tex, float u, float v, float3 p) {
    for (int depth = 0; depth < 9; depth++) {
        if (tex < 0) return f3(0.f);
        const float4 t0 = ldg4(sc.texs + 2 * tex);
        const int type = (int)f2u(t0.x);
        if (type == RTW_TEX_CONSTANT) return f3(t0.y, t0.z, t0.w);
        const float4 t1 = ldg4(sc.texs + 2 * tex + 1);
        if (type == RTW_TEX_CHECKER) {
            // Q18: sine formula literal, children dispatched by texture index
            float sines = det_sinf(10.f * p.x) * det_sinf(10.f - p.y) * det_sinf(10.f * p.z);
            tex = (sines < 0) ? (int)f2u(t1.x) : (int)f2u(t1.y);
            continue;
        }
        if (type == RTW_TEX_IMAGE) return image_lookup(sc, (int)f2u(t1.w), u, v);
        if (type == RTW_TEX_NOISE) {
            const float scale = t1.z;
            return f3(1.f) * 0.5f * (1.0f + det_sinf(scale * p.z + 5.f * perlin_turb(sc, (int)f2u(t1.w), scale * p)));
        }
        return f3(0.f, 0.f, 0.f);  // null texture
    }
    return f3(0.f);
}

}  // namespace rtw
