// material.h — sampling helpers, material callables and the rect-light sampler.
//   lib/sampling.cuh:15-60, lib/onb.cuh:9-35
//   material/lambertianMaterial.cu:29-81, metalMaterial.cu:32-64,
//   dielectricMaterial.cu:20-114, diffuseLight.cu:48-69, isotropicMaterial.cu:30-51,
//   normalMaterial.cu:21-31
//   pdf/rectPdf.cu:124-193 via pdf/mixturePdf.cu:25-38 (Q26)
#pragma once
#include "detmath.h"
#include "intersect.h"
#include "rng.h"
#include "texture.h"

namespace rtw {

// lib/sampling.cuh:15-22
RTW_HD float3 random_in_unit_disk(uint32_t& seed) {
    float a = rnd(seed) * 2.0f * RTW_PI;
    float sa, ca;
    det_sincosf(a, sa, ca);
    float3 xy = f3(sa, ca, 0);
    xy = xy * sqrtf(rnd(seed));
    return xy;
}

// lib/sampling.cuh:25-34; draws sequenced x, y, z
RTW_HD float3 randomInUnitSphere(uint32_t& seed) {
    float3 p;
    do {
        float x = rnd(seed);
        float y = rnd(seed);
        float z = rnd(seed);
        p = 2.0f * f3(x, y, z) - f3(1.0f, 1.0f, 1.0f);
    } while (dot(p, p) >= 1.0f);
    return p;
}

// lib/sampling.cuh:49-60 (Q10: factor 2 literal)
RTW_HD float3 randomCosineDirection(uint32_t& seed) {
    float r1 = rnd(seed);
    float r2 = rnd(seed);
    float phi = 2.f * RTW_PI * r1;
    float sphi, cphi;
    det_sincosf(phi, sphi, cphi);
    float x = cphi * 2 * sqrtf(r2);
    float y = sphi * 2 * sqrtf(r2);
    float z = sqrtf(1 - r2);
    return f3(x, y, z);
}

// lib/onb.cuh:9-35
struct Onb {
    float3 u, v, w;
    RTW_HD void buildFromW(float3 n) {
        w = normalize(n);
        float3 a;
        if ((w.x > 0.9f) || (w.x < -0.9f)) a = f3(0.f, 1.f, 0.f);
        else a = f3(1.f, 0.f, 0.f);
        v = normalize(cross(w, a));
        u = cross(w, v);
    }
    RTW_HD float3 local(float3 a) const { return a.x * u + a.y * v + a.z * w; }
};

// Result of a material callable: the PerRayData fields the path loop consumes.
#define RTW_EV_HIT    1   // Ray_Hit: path continues
#define RTW_EV_CANCEL 3   // Ray_Cancel / Ray_Finish / Ray_Miss: path ends
struct Scatter {
    int event;
    float3 origin;       // scattered_origin
    float3 direction;    // scattered_direction
    float3 attenuation;
    float3 radiance;     // emitted (light) or 0
};

// material/lambertianMaterial.cu:41-71 (Q14)
RTW_HD void mat_lambertian(const DScene& sc, int tex, const HitAttr& h, uint32_t& seed, Scatter& s) {
    Onb uvw;
    uvw.buildFromW(h.normal);
    float3 w = randomCosineDirection(seed);
    float pdf = w.z * RTW_INV_PI;
    float3 scatterDirection = normalize(uvw.local(w));
    s.event = RTW_EV_HIT;
    s.origin = h.point;
    s.direction = scatterDirection;
    s.radiance = f3(0.f);
    // scatteringPdf, lambertianMaterial.cu:29-38
    float cosine = dot(h.normal, normalize(scatterDirection));
    float scattered_pdf = (cosine < 0.f) ? 0.f : cosine / RTW_PI;
    if (scattered_pdf <= 0 || pdf <= 0) {
        s.event = RTW_EV_CANCEL;
        s.attenuation = f3(0.f);
        return;
    }
    s.attenuation = texture_eval(sc, tex, h.u, h.v, h.point);
}

// material/metalMaterial.cu:32-64 (Q12)
RTW_HD void mat_metal(const DScene& sc, int tex, float fuzz, float3 direction, const HitAttr& h, uint32_t& seed, Scatter& s) {
    float3 reflected = reflect(direction, h.normal);
    float3 scatterDirection = reflected + fuzz * randomInUnitSphere(seed);
    s.origin = h.point;
    s.direction = normalize(scatterDirection);
    s.radiance = f3(0.f);
    s.attenuation = texture_eval(sc, tex, h.u, h.v, h.point);
    s.event = (dot(s.direction, h.normal) <= 0.0f) ? RTW_EV_CANCEL : RTW_EV_HIT;
}

// material/dielectricMaterial.cu:20-26
RTW_HD float fresnelSchlick(float cosThetaI, float etaI, float etaT) {
    float r0 = (etaI - etaT) / (etaI + etaT);
    r0 = r0 * r0;
    return r0 + (1.f - r0) * det_pow5f((1.f - cosThetaI));
}

// material/dielectricMaterial.cu:37-114 (Q13); lightDirection == -direction (raygen.cu:39)
RTW_HD void mat_dielectric(float eta, float3 direction, const HitAttr& h, uint32_t& seed, Scatter& s) {
    float3 lightDirection = -direction;
    float3 unitDirection = normalize(-lightDirection);
    float3 localNormal;
    float etaI, etaT;
    if (dot(direction, h.normal) < 0.0f) {
        localNormal = h.normal;
        etaI = 1.0f;
        etaT = eta;
    } else {
        localNormal = -h.normal;
        etaI = eta;
        etaT = 1.0f;
    }
    float3 scatterDirection;
    float cosThetaI = fminf(dot(-unitDirection, localNormal), 1.0f);
    float sinThetaI = sqrtf(1.0f - cosThetaI * cosThetaI);
    if (((etaI / etaT) * sinThetaI) > 1.0f) {
        scatterDirection = reflect(unitDirection, localNormal);
    } else {
        float reflectProb = fresnelSchlick(cosThetaI, etaI, etaT);
        if (rnd(seed) < reflectProb) {
            scatterDirection = reflect(unitDirection, localNormal);
        } else {
            float sinThetaT = fminf((etaI / etaT) * sinThetaI, 1.0f);
            float cosThetaT = sqrtf(1.0f - sinThetaT * sinThetaT);
            scatterDirection = (etaI / etaT) * (unitDirection + cosThetaI * localNormal) - cosThetaT * localNormal;
        }
    }
    s.event = RTW_EV_HIT;
    s.origin = h.point;
    s.direction = scatterDirection;
    s.attenuation = f3(1.0f, 1.0f, 1.0f);
    s.radiance = f3(0.f);
}

// material/isotropicMaterial.cu:30-51 (Q15)
RTW_HD void mat_isotropic(const DScene& sc, int tex, const HitAttr& h, uint32_t& seed, Scatter& s) {
    float3 scatterDirection = randomInUnitSphere(seed);
    s.event = RTW_EV_HIT;
    s.origin = h.point;
    s.direction = scatterDirection;
    s.radiance = f3(0.f);
    s.attenuation = texture_eval(sc, tex, h.u, h.v, h.point);
}

// material/diffuseLight.cu:48-69 (Q17)
RTW_HD float3 mat_diffuse_light(const DScene& sc, int tex, float3 direction, const HitAttr& h) {
    if (dot(h.normal, direction) < 0.f) return texture_eval(sc, tex, h.u, h.v, h.point);
    return f3(0.f);
}

// miss/miss.cu:8-21 (Q27)
RTW_HD float3 miss_color(const DScene& sc, float3 direction) {
    if (sc.sky_light) {
        float3 unitDirection = normalize(direction);
        float t = 0.5f * (unitDirection.y + 1.0f);
        return (1.0f - t) * f3(1.0f, 1.0f, 1.0f) + t * f3(0.5f, 0.7f, 1.0f);
    }
    return f3(0.0f);
}

struct LightSample {   // lib/raydata.cuh:51-58
    float distance;
    float3 direction;
    float3 emission;
    float pdf;
};

// rect_{x,y,z}_generate, pdf/rectPdf.cu:124-193; draws sequenced a-axis, b-axis
RTW_HD LightSample rect_generate(const DScene& sc, int light_index, float3 origin, uint32_t& seed) {
    float ra = rnd(seed);
    float rb = rnd(seed);
    float pa = sc.rect_a0 + ra * (sc.rect_a1 - sc.rect_a0);
    float pb = sc.rect_b0 + rb * (sc.rect_b1 - sc.rect_b0);
    float3 random_point;
    if (sc.rect_axis == 0) random_point = f3(sc.rect_k, pa, pb);
    else if (sc.rect_axis == 1) random_point = f3(pa, sc.rect_k, pb);
    else random_point = f3(pa, pb, sc.rect_k);
    LightSample ls;
    ls.direction = random_point - origin;
    ls.distance = length(ls.direction);
    ls.pdf = 0.f;
    ls.emission = f3(0.f);
    if (ls.distance > 1.0e-6f) {
        ls.direction = ls.direction / ls.distance;
        const DLight& light = sc.lights[light_index];
        const float costa = dot(-ls.direction, light.normal);
        if (costa > 1.0e-6f) {
            ls.emission = light.emission * (float)sc.n_lights;
            ls.pdf = ls.distance * ls.distance / (light.area * costa);
        }
    }
    return ls;
}

// lib/raydata.cuh:166-170
RTW_HD float powerHeuristic(float a, float b) {
    const float t = a * a;
    return t / (t + b * b);
}

}  // namespace rtw
