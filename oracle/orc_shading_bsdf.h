// ORACLE — TEST INFRASTRUCTURE ONLY (parity pinned against oracle/_ref; see orc_types.h header).
// BSDFs + light sampling + blue noise: src/bsdf.cuh, src/light.cuh, src/blueNoiseRandGen.h:119-146
#pragma once
#include "orc_shading.h"

namespace orc {

// ---- src/bsdf.cuh:9-46 ----
static inline F2 ConcentricSampleDisk(F2 u) {
    // `2.0 * u - 1.0`: Float2 has only float operators, so the double literals narrow to float
    F2 uOffset = F2{2.0f * u.x - 1.0f, 2.0f * u.y - 1.0f};
    if (fabsf(uOffset.x) < 1e-10f && fabsf(uOffset.y) < 1e-10f) return F2{0, 0};
    float theta, r;
    if (fabsf(uOffset.x) > fabsf(uOffset.y)) { r = uOffset.x; theta = ORC_PI_OVER_4 * (uOffset.y / uOffset.x); }
    else { r = uOffset.y; theta = ORC_PI_OVER_2 - ORC_PI_OVER_4 * (uOffset.x / uOffset.y); }
    return F2{cosf(theta) * r, sinf(theta) * r};
}
static inline F2 UniformSampleDisk(F2 u) {
    float r = sqrtf(u.x);
    float theta = ORC_TWO_PI * u.y;
    return F2{r * cosf(theta), r * sinf(theta)};
}
static inline float max1f(float a, float b) { return (a < b) ? b : a; }       // src/linearMath.h:247
// CosineSampleHemisphere with DISK_SAMPLE_UNIFORM (src/bsdf.cuh:42-56)
static inline F3 CosineSampleHemisphere(F2 u) {
    F2 d = UniformSampleDisk(u);
    float z = sqrtf(max1f(0.0f, 1.0f - d.x * d.x - d.y * d.y));
    return F3{d.x, z, d.y};
}
// src/bsdf.cuh:71-101
static inline void LambertianSample(F2 r, F3& wo, F3 n) {
    F3 s = CosineSampleHemisphere(r);
    F3 u, v;
    LocalizeSample(n, u, v);
    wo = normalize(s.x * u + s.z * v + s.y * n);
}
static inline F3 LambertianBsdf(F3 albedo) { return albedo / ORC_M_PI; }
static inline float LambertianPdf(F3 wi, F3 n) { return rmax(dot(wi, n), SAFE_COSINE_EPSI) / ORC_M_PI; }

// src/bsdf.cuh:107-166
static inline float FresnelDialetric(float etaI, float etaT, float cosThetaI, float cosThetaT) {
    float R1 = etaT * cosThetaI, R2 = etaI * cosThetaT, R3 = etaI * cosThetaI, R4 = etaT * cosThetaT;
    float Rparl = (R1 - R2) / (R1 + R2);
    float Rperp = (R3 - R4) / (R3 + R4);
    return (float)((double)(Rparl * Rparl + Rperp * Rperp) / 2.0);
}
static inline float pow5(float e) { float e2 = e * e; return e2 * e2 * e; }
static inline F3 FresnelShlick(F3 F0, float cosTheta) { return F0 + (f3(1.0f) - F0) * pow5(1.0f - cosTheta); }

static inline void PerfectReflectionRefraction(float etaI, float etaT, bool into, F3 normal, float nDotRay, float rnd, F3 rayDir,
                                               F3& next, float& rayoffset) {
    if (!into) { float t = etaI; etaI = etaT; etaT = t; }
    const float eta = etaI / etaT;
    float cosThetaI = -nDotRay;
    float sin2ThetaI = max1f(0, (float)(1.0 - (double)(cosThetaI * cosThetaI)));
    float sin2ThetaT = eta * eta * sin2ThetaI;
    float cosThetaT = sqrtf(max1f(0, (float)(1.0 - (double)sin2ThetaT)));   // sqrt(float) -> float overload in device code
    if (sin2ThetaT >= 1.0) {
        next = rayDir - normal * nDotRay * 2.0f;
    } else {
        float fresnel = FresnelDialetric(etaI, etaT, cosThetaI, cosThetaT);
        if (rnd < fresnel) next = rayDir - normal * nDotRay * 2.0f;
        else { next = eta * rayDir + (eta * cosThetaI - cosThetaT) * normal; rayoffset = (float)((double)rayoffset * -1.0); }
    }
    next = normalize(next);
}

static inline F3 reflect3f(F3 i, F3 n) { return i - 2.0f * n * dot(n, i); }            // src/linearMath.h:489

// GGX terms shared by MacrofacetReflectionSample / MacrofacetReflection (src/bsdf.cuh:239-298)
static inline void ggx_eval(F3 wn, F3 wo, F3 wi, F3 wh, F3 F0, F3 albedo, float alpha2, F3& brdfOverPdf, F3& brdf, float& pdf) {
    float cosThetaWoWh = rmax(SAFE_COSINE_EPSI, dot(wh, wo));
    F3 F = FresnelShlick(F0, cosThetaWoWh);
    float cosThetaWo = clampf(dot(wo, wn), SAFE_COSINE_EPSI, 1.0f - SAFE_COSINE_EPSI);
    float cosThetaWi = rmax(SAFE_COSINE_EPSI, dot(wi, wn));
    float tanThetaWo = sqrtf(1.0f - cosThetaWo * cosThetaWo) / cosThetaWo;
    float G = 1.0f / (1.0f + (sqrtf(1.0f + alpha2 * tanThetaWo * tanThetaWo) - 1.0f) / 2.0f);
    float cosThetaWh = rmax(SAFE_COSINE_EPSI, dot(wh, wn));
    float cosTheta2Wh = cosThetaWh * cosThetaWh;
    float tanTheta2Wh = (1.0f - cosTheta2Wh) / cosTheta2Wh;
    float e = tanTheta2Wh / alpha2 + 1.0f;
    float D = 1.0f / (ORC_M_PI * (alpha2 * cosTheta2Wh * cosTheta2Wh) * (e * e));
    brdf = (albedo * F) * (D * G) / (4.0f * cosThetaWo * cosThetaWi);
    pdf = (D * cosThetaWh) / (4.0f * cosThetaWoWh);
    brdfOverPdf = (albedo * F) * (G * cosThetaWoWh) / (cosThetaWh * cosThetaWo);
}

static inline F3 ggx_sample_normal(F2 r, float alpha2, F3 normal) {
    // `sqrt`, `cos`, `sin` on float arguments resolve to the float overloads in device code
    float cosTheta = 1.0f / sqrtf(1.0f + alpha2 * r.x / (1.0f - r.x));
    float sinTheta = sqrtf(1.0f - cosTheta * cosTheta);
    float phi = ORC_TWO_PI * r.y;
    F3 local = F3{sinTheta * cosf(phi), cosTheta, sinTheta * sinf(phi)};
    F3 t, b;
    LocalizeSample(normal, t, b);
    return normalize(local.x * t + local.z * b + local.y * normal);
}
// src/bsdf.cuh:168-263
static inline void MacrofacetReflectionSample(F2 r, F2 r2, F3 raydir, F3& nextdir, F3 normal, F3 surfaceNormal,
                                              F3& brdfOverPdf, F3& brdf, float& pdf, F3 F0, F3 albedo, float alpha) {
    float alpha2 = alpha * alpha;
    F3 sampledNormal = ggx_sample_normal(r, alpha2, normal);
    nextdir = normalize(reflect3f(raydir, sampledNormal));
    if (dot(nextdir, surfaceNormal) < 0) {
        F3 sn2 = ggx_sample_normal(r2, alpha2, normal);      // inner-scope sampledNormal: wh below keeps the FIRST one
        nextdir = normalize(reflect3f(raydir, sn2));
        if (dot(nextdir, surfaceNormal) < 0) nextdir = normalize(reflect3f(raydir, normal));
    }
    ggx_eval(normal, -raydir, nextdir, sampledNormal, F0, albedo, alpha2, brdfOverPdf, brdf, pdf);
}
// src/bsdf.cuh:265-298 (the early-out assignments are overwritten: there is no return after them)
static inline void MacrofacetReflection(F3& brdfOverPdf, F3& brdf, float& pdf, F3 wn, F3 wo, F3 wi, F3 F0, F3 albedo, float alpha) {
    float alpha2 = alpha * alpha;
    F3 wh = normalize(wi + wo);
    ggx_eval(wn, wo, wi, wh, F0, albedo, alpha2, brdfOverPdf, brdf, pdf);
}

// ---- blue noise: src/blueNoiseRandGen.h:119-146 ----
struct BlueNoise { const uint8_t* sobol; const uint8_t* scramble; const uint8_t* rank; };
static inline float blue_rand(const BlueNoise& bn, int px, int py, int sampleIndex, int dim) {
    px &= 127; py &= 127; sampleIndex &= 255;
    int ranked = sampleIndex ^ bn.rank[dim + (px + py * 128) * 8];
    int value = bn.sobol[dim + ranked * 256];
    value = value ^ bn.scramble[(dim % 8) + (px + py * 128) * 8];
    return (value + 0.5f) / 256.0f;
}

// ---- src/light.cuh:9-31 ----
static inline int BinarySearch(const float* a, int left, int right, float target) {
    while (right - left > 1) {
        int mid = (left + right) / 2;
        if (a[mid] < target) left = mid; else right = mid;
    }
    return left;
}

struct Env {
    const float* sky; const float* skyCdf; const float* sun; const float* sunCdf;   // float4 maps + inclusive CDFs
    F3 sunDir; float sunAngleCosThetaMax;
};

// src/light.cuh:73-273 (RENDER_SPHERE_LIGHT 0): sky-vs-sun by flux, CDF inversion, equal-area maps
static inline void SampleLight(const Env& e, F3& dir, float& pdf, int& lightIdx, const F4& rnd) {
    const float maxSkyCdf = e.skyCdf[SKY_SIZE - 1];
    const float maxSunCdf = e.sunCdf[SUN_SIZE - 1];
    const float totalSkyLum = maxSkyCdf * ORC_TWO_PI / SKY_SIZE;
    const float totalSunLum = maxSunCdf * ORC_TWO_PI * (1.0f - e.sunAngleCosThetaMax) / SUN_SIZE;
    const float chooseSky = totalSkyLum / (totalSkyLum + totalSunLum);
    if (chooseSky > rnd.y) {
        const int idx = BinarySearch(e.skyCdf, 0, SKY_SIZE - 2, rnd.x * maxSkyCdf) + 1;
        float p = (e.skyCdf[idx] - e.skyCdf[idx - 1]) / maxSkyCdf;
        p = p * SKY_SIZE / ORC_TWO_PI;
        float u = ((idx % SKY_WIDTH) + 0.5f) / SKY_WIDTH;
        float v = ((idx / SKY_WIDTH) + 0.5f) / SKY_HEIGHT;
        dir = EqualAreaMap(u, v);
        pdf = p * 1.0f * chooseSky;
    } else {
        const int idx = BinarySearch(e.sunCdf, 0, SUN_SIZE - 2, rnd.x * maxSunCdf) + 1;
        float p = (e.sunCdf[idx] - e.sunCdf[idx - 1]) / maxSunCdf;
        p = p * SUN_SIZE / (ORC_TWO_PI * (1.0f - e.sunAngleCosThetaMax));
        float u = ((idx % SUN_WIDTH) + 0.5f) / SUN_WIDTH;
        float v = ((idx / SUN_WIDTH) + 0.5f) / SUN_HEIGHT;
        dir = EqualAreaMapCone(e.sunDir, u, v, e.sunAngleCosThetaMax);
        pdf = p * 1.0f;                      // the (1 - chooseSky) factor is commented out in the reference (:224)
    }
    lightIdx = ENV_LIGHT_ID;
}

static inline F3 smoothstep3f(F3 a, F3 b, float w) { return a + (w * w * (3.0f - 2.0f * w)) * (b - a); }    // linearMath.h:495

// src/light.cuh:33-71
static inline F3 EnvLight2(const Env& e, F3 raydir) {
    F3 color = f3(0.0f);
    {
        F3 skyColor = sample_env(e.sky, SKY_WIDTH, SKY_HEIGHT, EqualAreaMapDir(raydir), true);
        F3 mist = f3(0.2f);
        float blend = clampf((raydir.y + 0.4f) * (1.0f / 0.5f));
        color = color + smoothstep3f(mist, skyColor, blend);
    }
    {
        F2 uv;
        if (EqualAreaMapConeInv(uv, e.sunDir, raydir, e.sunAngleCosThetaMax))
            color = color + sample_env(e.sun, SUN_WIDTH, SUN_HEIGHT, uv, false);
    }
    return color;
}

}  // namespace orc
