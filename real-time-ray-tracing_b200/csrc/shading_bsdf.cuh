// BSDFs and light sampling of the path tracer (device).  See shading.cuh for the citation map.
#pragma once
#include "shading.cuh"

namespace rtrt {

// ---- src/bsdf.cuh:9-46 ----
VD Float2 concentric_sample_disk(Float2 u) {
    const Float2 o{2.0f * u.x - 1.0f, 2.0f * u.y - 1.0f};
    if (fabsf(o.x) < 1e-10f && fabsf(o.y) < 1e-10f) return Float2{0, 0};
    float theta, r;
    if (fabsf(o.x) > fabsf(o.y)) { r = o.x; theta = RTRT_PI_OVER_4 * (o.y / o.x); }
    else { r = o.y; theta = RTRT_PI_OVER_2 - RTRT_PI_OVER_4 * (o.x / o.y); }
    return Float2{cosf(theta) * r, sinf(theta) * r};
}
VD Float2 uniform_sample_disk(Float2 u) {
    const float r = sqrtf(u.x);
    const float theta = RTRT_TWO_PI * u.y;
    return Float2{r * cosf(theta), r * sinf(theta)};
}
// CosineSampleHemisphere, DISK_SAMPLE_UNIFORM (src/bsdf.cuh:42-56)
VD Float3 cosine_sample_hemisphere(Float2 u) {
    const Float2 d = uniform_sample_disk(u);
    const float z = sqrtf(max1f(0.0f, 1.0f - d.x * d.x - d.y * d.y));
    return Float3{d.x, z, d.y};
}
// src/bsdf.cuh:71-101
VD Float3 lambertian_sample(Float2 r, Float3 n) {
    const Float3 s = cosine_sample_hemisphere(r);
    Float3 u, v;
    localize_sample(n, u, v);
    return normalize(s.x * u + s.z * v + s.y * n);
}
VD Float3 lambertian_bsdf(Float3 albedo) { return albedo / RTRT_M_PI; }
VD float lambertian_pdf(Float3 wi, Float3 n) { return sel_max(dot(wi, n), RTRT_SAFE_COSINE_EPSI) / RTRT_M_PI; }

// src/bsdf.cuh:107-166
VD float fresnel_dielectric(float etaI, float etaT, float cosI, float cosT) {
    const float R1 = etaT * cosI, R2 = etaI * cosT, R3 = etaI * cosI, R4 = etaT * cosT;
    const float Rparl = (R1 - R2) / (R1 + R2);
    const float Rperp = (R3 - R4) / (R3 + R4);
    return (Rparl * Rparl + Rperp * Rperp) / 2.0f;
}
VD float pow5(float e) { const float e2 = e * e; return e2 * e2 * e; }
VD Float3 fresnel_schlick(Float3 F0, float cosTheta) { return F0 + (f3(1.0f) - F0) * pow5(1.0f - cosTheta); }

VD void perfect_reflection_refraction(float etaI, float etaT, bool into, Float3 normal, float nDotRay, float rnd, Float3 rayDir,
                                      Float3& next, float& rayoffset) {
    if (!into) { const float t = etaI; etaI = etaT; etaT = t; }
    const float eta = etaI / etaT;
    const float cosI = -nDotRay;
    const float sin2I = max1f(0.0f, (float)(1.0 - (double)(cosI * cosI)));
    const float sin2T = eta * eta * sin2I;
    const float cosT = sqrtf(max1f(0.0f, (float)(1.0 - (double)sin2T)));
    if (sin2T >= 1.0f) {
        next = rayDir - normal * nDotRay * 2.0f;
    } else {
        const float fr = fresnel_dielectric(etaI, etaT, cosI, cosT);
        if (rnd < fr) next = rayDir - normal * nDotRay * 2.0f;
        else { next = eta * rayDir + (eta * cosI - cosT) * normal; rayoffset = -rayoffset; }
    }
    next = normalize(next);
}

VD Float3 reflect3f(Float3 i, Float3 n) { return i - 2.0f * n * dot(n, i); }       // src/linearMath.h:489

// GGX evaluation shared by MacrofacetReflectionSample / MacrofacetReflection (src/bsdf.cuh:239-298)
VD void ggx_eval(Float3 wn, Float3 wo, Float3 wi, Float3 wh, Float3 F0, Float3 albedo, float alpha2, Float3& brdf, float& pdf) {
    const float cosWoWh = sel_max(RTRT_SAFE_COSINE_EPSI, dot(wh, wo));
    const Float3 F = fresnel_schlick(F0, cosWoWh);
    const float cosWo = clampf(dot(wo, wn), RTRT_SAFE_COSINE_EPSI, 1.0f - RTRT_SAFE_COSINE_EPSI);
    const float cosWi = sel_max(RTRT_SAFE_COSINE_EPSI, dot(wi, wn));
    const float tanWo = sqrtf(1.0f - cosWo * cosWo) / cosWo;
    const float G = 1.0f / (1.0f + (sqrtf(1.0f + alpha2 * tanWo * tanWo) - 1.0f) / 2.0f);
    const float cosWh = sel_max(RTRT_SAFE_COSINE_EPSI, dot(wh, wn));
    const float cos2Wh = cosWh * cosWh;
    const float tan2Wh = (1.0f - cos2Wh) / cos2Wh;
    const float e = tan2Wh / alpha2 + 1.0f;
    const float D = 1.0f / (RTRT_M_PI * (alpha2 * cos2Wh * cos2Wh) * (e * e));
    brdf = (albedo * F) * (D * G) / (4.0f * cosWo * cosWi);
    pdf = (D * cosWh) / (4.0f * cosWoWh);
}
VD Float3 ggx_sample_normal(Float2 r, float alpha2, Float3 normal) {
    const float cosTheta = 1.0f / sqrtf(1.0f + alpha2 * r.x / (1.0f - r.x));
    const float sinTheta = sqrtf(1.0f - cosTheta * cosTheta);
    const float phi = RTRT_TWO_PI * r.y;
    const Float3 local{sinTheta * cosf(phi), cosTheta, sinTheta * sinf(phi)};
    Float3 t, b;
    localize_sample(normal, t, b);
    return normalize(local.x * t + local.z * b + local.y * normal);
}
// src/bsdf.cuh:168-263
VD void microfacet_reflection_sample(Float2 r, Float2 r2, Float3 raydir, Float3& nextdir, Float3 normal, Float3 surfaceNormal,
                                     Float3& brdf, float& pdf, Float3 F0, Float3 albedo, float alpha) {
    const float alpha2 = alpha * alpha;
    const Float3 sampledNormal = ggx_sample_normal(r, alpha2, normal);
    nextdir = normalize(reflect3f(raydir, sampledNormal));
    if (dot(nextdir, surfaceNormal) < 0) {
        const Float3 sn2 = ggx_sample_normal(r2, alpha2, normal);      // shadowed variable in the reference: wh keeps the first normal
        nextdir = normalize(reflect3f(raydir, sn2));
        if (dot(nextdir, surfaceNormal) < 0) nextdir = normalize(reflect3f(raydir, normal));
    }
    ggx_eval(normal, -raydir, nextdir, sampledNormal, F0, albedo, alpha2, brdf, pdf);
}
// src/bsdf.cuh:265-298
VD void microfacet_reflection(Float3& brdf, float& pdf, Float3 wn, Float3 wo, Float3 wi, Float3 F0, Float3 albedo, float alpha) {
    ggx_eval(wn, wo, wi, normalize(wi + wo), F0, albedo, alpha * alpha, brdf, pdf);
}

// src/light.cuh:9-31
VD int binary_search(const float* __restrict__ a, int left, int right, float target) {
    while (right - left > 1) {
        const int mid = (left + right) / 2;
        if (__ldg(a + mid) < target) left = mid; else right = mid;
    }
    return left;
}

// src/light.cuh:73-273 with RENDER_SPHERE_LIGHT 0: sky-vs-sun choice by flux, CDF inversion, equal-area maps
VD void sample_light(const EnvMaps& e, Float3& dir, float& pdf, int& lightIdx, const Float4& rnd) {
    const float maxSkyCdf = __ldg(e.skyCdf + kSkySize - 1);
    const float maxSunCdf = __ldg(e.sunCdf + kSunSize - 1);
    const float totalSkyLum = maxSkyCdf * RTRT_TWO_PI / kSkySize;
    const float totalSunLum = maxSunCdf * RTRT_TWO_PI * (1.0f - e.sunAngleCosThetaMax) / kSunSize;
    const float chooseSky = totalSkyLum / (totalSkyLum + totalSunLum);
    if (chooseSky > rnd.y) {
        const int idx = binary_search(e.skyCdf, 0, kSkySize - 2, rnd.x * maxSkyCdf) + 1;
        float p = (__ldg(e.skyCdf + idx) - __ldg(e.skyCdf + idx - 1)) / maxSkyCdf;
        p = p * kSkySize / RTRT_TWO_PI;
        const float u = ((idx % kSkyW) + 0.5f) / kSkyW;
        const float v = ((idx / kSkyW) + 0.5f) / kSkyH;
        dir = equal_area_map(u, v);
        pdf = p * 1.0f * chooseSky;
    } else {
        const int idx = binary_search(e.sunCdf, 0, kSunSize - 2, rnd.x * maxSunCdf) + 1;
        float p = (__ldg(e.sunCdf + idx) - __ldg(e.sunCdf + idx - 1)) / maxSunCdf;
        p = p * kSunSize / (RTRT_TWO_PI * (1.0f - e.sunAngleCosThetaMax));
        const float u = ((idx % kSunW) + 0.5f) / kSunW;
        const float v = ((idx / kSunW) + 0.5f) / kSunH;
        dir = equal_area_map_cone(e.sunDir, u, v, e.sunAngleCosThetaMax);
        pdf = p * 1.0f;                       // the (1 - chooseSky) factor is commented out in the reference (:224)
    }
    lightIdx = kEnvLightId;
}

VD Float3 smoothstep3f(Float3 a, Float3 b, float w) { return a + (w * w * (3.0f - 2.0f * w)) * (b - a); }   // linearMath.h:495

// src/light.cuh:33-71
VD Float3 env_light(const EnvMaps& e, Float3 raydir) {
    Float3 color = f3(0.0f);
    {
        const Float3 skyColor = sample_env(e.sky, kSkyW, kSkyH, equal_area_map_dir(raydir), true);
        const float blend = clampf((raydir.y + 0.4f) * (1.0f / 0.5f));
        color = color + smoothstep3f(f3(0.2f), skyColor, blend);
    }
    {
        Float2 uv;
        if (equal_area_map_cone_inv(uv, e.sunDir, raydir, e.sunAngleCosThetaMax))
            color = color + sample_env(e.sun, kSunW, kSunH, uv, false);
    }
    return color;
}

}  // namespace rtrt
