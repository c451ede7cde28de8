// ORACLE — TEST INFRASTRUCTURE ONLY (parity pinned against oracle/_ref; see orc_types.h header).
// Stage 2: the path-trace megakernel, one pixel at a time on the CPU, restating
//   src/pathtrace.cuh:11-128         PathTrace
//   src/raygen.cuh:7-63              GenerateRay, GetRayConeWidth
//   src/traverse.cuh:9-225           UpdateMaterial, RaySceneIntersect
//   src/surfaceInteraction.cuh       GlossySurfaceInteraction, DiffuseSurfaceInteraction
//   src/light.cuh:275-305            GetLightSource
#include "orc_shading_bsdf.h"
#include <thread>
#include <vector>

namespace orc {

struct RayState {            // the live subset of src/kernel.cuh:233-284
    F3 orig, dir, beta0, beta1, pos, normal, fakeNormal, albedo, centerRaydir;
    F2 uv;
    int matId, matType, lightIdx, objectIdx;
    float offset, normalDotRayDir, depth, rayConeWidth, rayConeSpread;
    bool isRayIntoSurface, hitLight, hit, isDiffuseRay, isDiffuse, isHitProcessed, isOccluded, isShadowRay;
};

struct PtScene {
    Scene bvh;
    const SurfaceMaterial* materials; const int* materialsIdx; int numMaterials, numMaterialsIdx;
    Camera camera; HistoryCamera historyCamera;
    Env env;
    BlueNoise bn;
    Mipmap tex[2];
    int frameNum;
};

// src/raygen.cuh:45-63
static float GetRayConeWidth(const Camera& cam, int ix, int iy) {
    F2 pixelCenter = (F2{(float)ix, (float)iy} + 0.5f) - F2{cam.resolution.x, cam.resolution.y} / 2;
    F2 pixelOffset = F2{copysignf(0.5f, pixelCenter.x), copysignf(0.5f, pixelCenter.y)};
    F2 uvNear = (pixelCenter - pixelOffset) * cam.inversedResolution * 2;
    F2 uvFar = (pixelCenter + pixelOffset) * cam.inversedResolution * 2;
    F2 halfFovLength = F2{tanf(cam.fov.x / 2), tanf(cam.fov.y / 2)};
    F2 nearP = uvNear * halfFovLength, farP = uvFar * halfFovLength;
    float angleNear = atanf(sqrtf(nearP.x * nearP.x + nearP.y * nearP.y));
    float angleFar = atanf(sqrtf(farP.x * farP.x + farP.y * farP.y));
    return angleFar - angleNear;
}

// src/raygen.cuh:7-38
static void GenerateRay(F3& orig, F3& dir, F3& centerDir, F2& sampleUv, const Camera& cam, int ix, int iy, F2 rndPixel, F2 rndAperture) {
    F2 uv = (F2{(float)ix, (float)iy} + rndPixel) * cam.inversedResolution;
    F2 uvCenter = (F2{(float)ix, (float)iy} + 0.5f) * cam.inversedResolution;
    sampleUv = uv;
    uv = uv * -2.0f + 1.0f;
    uvCenter = uvCenter * -2.0f + 1.0f;
    F3 p = cam.adjustedFront + cam.adjustedLeft * uv.x + cam.adjustedUp * uv.y;
    F3 pc = cam.adjustedFront + cam.adjustedLeft * uvCenter.x + cam.adjustedUp * uvCenter.y;
    F2 disk = ConcentricSampleDisk(rndAperture);
    F3 pa = disk.x * cam.apertureLeft + disk.y * cam.apertureUp;
    orig = cam.pos + pa;
    dir = normalize(p - pa);
    centerDir = normalize(pc);
}

// src/traverse.cuh:9-56
static void UpdateMaterial(const PtScene& s, RayState& r) {
    if (!r.hit) { r.matType = MAT_SKY; r.matId = 99999; }
    else {
        r.matId = (r.objectIdx >= 0 && r.objectIdx < s.numMaterialsIdx) ? s.materialsIdx[r.objectIdx] : 6;
        SurfaceMaterial def{f3(0.8f), PERFECT_REFLECTION, F3{0.56f, 0.57f, 0.58f}, 0.05f};
        SurfaceMaterial mat = (r.matId >= 0 && r.matId < s.numMaterials) ? s.materials[r.matId] : def;
        r.matType = (int)mat.type;
    }
    if (r.isShadowRay) {
        if ((r.matType == EMISSIVE && r.lightIdx == r.objectIdx) || (r.matType == MAT_SKY && r.lightIdx == ENV_LIGHT_ID)) r.hitLight = true;
        else r.isOccluded = true;
    } else {
        r.hitLight = r.matType == EMISSIVE || r.matType == MAT_SKY;
    }
    r.isDiffuse = (r.matType == LAMBERTIAN_DIFFUSE) || (r.matType == MICROFACET_REFLECTION);
}

// src/traverse.cuh:64-225
static void RaySceneIntersect(const PtScene& s, RayState& r, unsigned long long& rays, unsigned long long& nodeVisits, unsigned long long& leafTests) {
    if (r.hitLight || !r.isHitProcessed || r.isOccluded) return;
    r.isHitProcessed = false;
    Ray ray{r.orig, 0, r.dir, 0};
    F3 inv = SafeDivide3f(f3(1.0f), ray.dir);
    HitRecord h = init_hit();
    h.fakeNormal = r.fakeNormal;                       // not reset by the reference (reference member, :77)
    TraverseBvh(s.bvh, ray, inv, h, 0);
    rays++; nodeVisits += h.nodeVisits; leafTests += h.leafTests;
    r.pos = h.point; r.normal = h.normal; r.fakeNormal = h.fakeNormal; r.objectIdx = h.objectIdx; r.offset = h.rayOffset; r.uv = h.uv;
    r.normalDotRayDir = dot(r.normal, ray.dir);
    r.isRayIntoSurface = r.normalDotRayDir < 0;
    if (!r.isRayIntoSurface) { r.normal = -r.normal; r.normalDotRayDir = -r.normalDotRayDir; }
    if (dot(r.fakeNormal, r.normal) < 0) r.fakeNormal = -r.fakeNormal;
    r.hit = (h.t < RayMax);
    r.depth = h.t;
    if (!r.hit) { r.normal = f3(0, -1, 0); r.fakeNormal = f3(0, -1, 0); }
    else r.rayConeWidth += r.rayConeSpread * h.t;
    UpdateMaterial(s, r);
}

// src/surfaceInteraction.cuh:11-34
static void GlossySurfaceInteraction(RayState& r, float rnd) {
    if (r.hitLight || r.isDiffuse || r.isOccluded) return;
    r.isHitProcessed = true;
    if (r.matType == PERFECT_REFLECTION) {
        r.dir = normalize(r.dir - r.normal * dot(r.dir, r.normal) * 2.0f);
        r.orig = r.pos + r.offset * r.normal;
    } else if (r.matType == PERFECT_FRESNEL_REFLECTION_REFRACTION) {
        F3 next; float off = r.offset;
        PerfectReflectionRefraction(1.0f, 1.33f, r.isRayIntoSurface, r.normal, r.normalDotRayDir, rnd, r.dir, next, off);
        r.dir = next;
        r.orig = r.pos + off * r.normal;
    }
}

static inline F3 pow3f(F3 v, float a) { return F3{powf(v.x, a), powf(v.y, a), powf(v.z, a)}; }
static inline F3 xyz(F4 v) { return F3{v.x, v.y, v.z}; }

// one triplanar projection of src/surfaceInteraction.cuh:96-155
static void triplanar_axis(const PtScene& s, F2 uv, float lod, F3 normal, F3 w0, bool useAlt, F3 wAlt, F3& albedo, float& ao, F3& nrm, float& rough) {
    uv = uv * 0.5f;                                   // uvScale
    F4 c = sample_tex_lod(s.tex[0], uv, lod);
    albedo = pow3f(xyz(c), 2.2f);
    ao = c.w;
    F4 c2 = sample_tex_lod(s.tex[1], uv, lod);
    F3 n = xyz(c2) - 0.5f;
    rough = c2.w;
    F3 w = useAlt ? wAlt : w0;
    F3 u = cross(normal, w);
    F3 v = cross(normal, u);
    nrm = normalize(u * n.x + v * n.y + normal * n.z);
}

// src/surfaceInteraction.cuh:36-310
static void DiffuseSurfaceInteraction(int bounce, const PtScene& s, RayState& r, F3& beta, const F4& rnd, const F4& rnd2) {
    if (r.hitLight || !r.isDiffuse || r.isOccluded) return;
    r.isDiffuseRay = true;
    r.lightIdx = DEFAULT_LIGHT_ID;
    r.isHitProcessed = true;
    SurfaceMaterial mat = s.materials[r.matId];
    F3 normal = r.fakeNormal, surfaceNormal = r.normal;
    F3 albedo; float ao, roughness;
    {
        const float uvScale = 0.5f;
        float tw = (float)s.tex[0].w[0], th = (float)s.tex[0].h[0];
        float lod = log2f(r.rayConeWidth * uvScale * sqrtf(tw * tw + th * th));
        F3 aX, aY, aZ, nX, nY, nZ; float oX, oY, oZ, rX, rY, rZ;
        triplanar_axis(s, F2{r.pos.y, r.pos.z}, lod, normal, f3(0, 1, 0), fabsf(normal.y) > 0.999f, f3(0, 0, 1), aX, oX, nX, rX);
        triplanar_axis(s, F2{r.pos.x, r.pos.z}, lod, normal, f3(1, 0, 0), fabsf(normal.x) > 0.999f, f3(0, 0, 1), aY, oY, nY, rY);
        triplanar_axis(s, F2{r.pos.x, r.pos.y}, lod, normal, f3(0, 1, 0), fabsf(normal.y) > 0.999f, f3(1, 0, 0), aZ, oZ, nZ, rZ);
        float wx = surfaceNormal.x * surfaceNormal.x, wy = surfaceNormal.y * surfaceNormal.y, wz = surfaceNormal.z * surfaceNormal.z;
        albedo = aX * wx + aY * wy + aZ * wz;
        ao = oX * wx + oY * wy + oZ * wz;
        F3 texNormal = normalize(nX * wx + nY * wy + nZ * wz);
        roughness = rX * wx + rY * wy + rZ * wz;
        (void)ao; (void)roughness;
        normal = texNormal;
        r.fakeNormal = normal;
    }
    if (bounce == 0) r.albedo = albedo * (1.0f + fabsf(dot(normal, r.centerRaydir)));
    F3 rayDir = r.dir;
    F3 lightDir; float lightPdf = 1; int lightIdx;
    SampleLight(s.env, lightDir, lightPdf, lightIdx, rnd2);
    F3 surfDir = f3(0.0f);
    F3 sBop = f3(0.0f), sBsdf = f3(0.0f); float sPdf = 0;
    F3 lBop = f3(0.0f), lBsdf = f3(0.0f); float lPdf = 0;
    F2 r2{rnd.x, rnd.y}, r22{rnd.z, rnd.w};
    if (r.matType == LAMBERTIAN_DIFFUSE) {
        LambertianSample(r2, surfDir, normal);
        sBop = albedo; sBsdf = LambertianBsdf(albedo); sPdf = LambertianPdf(surfDir, normal);
        lBop = albedo; lBsdf = LambertianBsdf(albedo); lPdf = LambertianPdf(lightDir, normal);
    } else if (r.matType == MICROFACET_REFLECTION) {
        MacrofacetReflectionSample(r2, r22, rayDir, surfDir, normal, surfaceNormal, sBop, sBsdf, sPdf, mat.F0, albedo, mat.alpha);
        MacrofacetReflection(lBop, lBsdf, lPdf, normal, -rayDir, lightDir, mat.F0, albedo, mat.alpha);
    }
    (void)sBop; (void)lBop; (void)lPdf;
    auto nanzero3 = [](F3& v) { if (v.x != v.x || v.y != v.y || v.z != v.z) v = f3(0.0f); };
    nanzero3(lBsdf); nanzero3(sBsdf);
    if (sPdf != sPdf) sPdf = 0;
    if (lightPdf != lightPdf) lightPdf = 0;
    float powerHeuristicSurface = (sPdf * sPdf) / (sPdf * sPdf + lightPdf * lightPdf);
    const float minFinalPdf = 1e-5f;
    float cosThetaWi;
    if (rnd.w < powerHeuristicSurface) {
        if (dot(r.normal, surfDir) < 0) { r.isOccluded = true; return; }
        cosThetaWi = rmax(SAFE_COSINE_EPSI, dot(surfDir, normal));
        beta = sBsdf * cosThetaWi / rmax(sPdf, minFinalPdf);
        r.dir = surfDir;
    } else {
        if (dot(r.normal, lightDir) < 0) { r.isOccluded = true; return; }
        cosThetaWi = rmax(SAFE_COSINE_EPSI, dot(lightDir, normal));
        beta = lBsdf * cosThetaWi / rmax(lightPdf, minFinalPdf);
        r.dir = lightDir;
        r.lightIdx = lightIdx;
        r.isShadowRay = true;
    }
    nanzero3(beta);
    r.orig = r.pos + r.offset * r.normal;
}

struct PixelOut { F3 color; uint16_t mask; F3 normal; F3 albedo; float depth; F2 motion; };

static PixelOut trace_pixel(const PtScene& s, int ix, int iy, unsigned long long* cnt) {
    RayState r{};
    r.beta0 = f3(1.0f); r.beta1 = f3(1.0f);
    r.isDiffuseRay = false; r.hitLight = false; r.lightIdx = DEFAULT_LIGHT_ID; r.isHitProcessed = true; r.isOccluded = false;
    r.isShadowRay = false; r.normal = f3(0, -1, 0); r.albedo = f3(1.0f); r.rayConeWidth = 0.0f;
    r.rayConeSpread = GetRayConeWidth(s.camera, ix, iy);
    F4 rnd[4];
    for (int k = 0; k < 4; ++k) {
        int si = s.frameNum * 4 + k;
        rnd[k] = F4{blue_rand(s.bn, ix, iy, si, 0), blue_rand(s.bn, ix, iy, si, 1), blue_rand(s.bn, ix, iy, si, 2), blue_rand(s.bn, ix, iy, si, 3)};
    }
    F2 sampleUv;
    GenerateRay(r.orig, r.dir, r.centerRaydir, sampleUv, s.camera, ix, iy, F2{rnd[0].x, rnd[0].y}, F2{rnd[0].z, rnd[0].w});
    RaySceneIntersect(s, r, cnt[0], cnt[1], cnt[2]);
    float outputDepth = r.depth;
    uint16_t mask = (uint16_t)r.matId;
    F2 motion{0, 0};
    if (r.hit) motion = WorldToScreenSpace(s.historyCamera, r.pos, s.camera.tanHalfFov) - sampleUv;
    motion = motion + 0.5f;
    GlossySurfaceInteraction(r, rnd[0].x); RaySceneIntersect(s, r, cnt[0], cnt[1], cnt[2]);
    GlossySurfaceInteraction(r, rnd[0].y); RaySceneIntersect(s, r, cnt[0], cnt[1], cnt[2]);
    GlossySurfaceInteraction(r, rnd[0].z);
    DiffuseSurfaceInteraction(0, s, r, r.beta1, rnd[0], rnd[1]);
    F3 outputNormal = r.fakeNormal;
    RaySceneIntersect(s, r, cnt[0], cnt[1], cnt[2]);
    GlossySurfaceInteraction(r, rnd[0].w);
    DiffuseSurfaceInteraction(1, s, r, r.beta0, rnd[2], rnd[3]);
    RaySceneIntersect(s, r, cnt[0], cnt[1], cnt[2]);
    F3 L0 = f3(0.0f);                                                     // GetLightSource src/light.cuh:275-305
    if (r.hitLight && !r.isOccluded && r.matType == MAT_SKY) L0 = EnvLight2(s.env, r.dir);
    F3 L1 = L0 * r.beta0;
    F3 L2 = L1 * r.beta1;
    auto nan3 = [](F3 v) { return v.x != v.x || v.y != v.y || v.z != v.z; };
    if (nan3(L2)) L2 = f3(0.0f);
    if (nan3(outputNormal)) outputNormal = f3(0.0f);
    if (outputDepth != outputDepth) outputDepth = 0;
    if (motion.x != motion.x || motion.y != motion.y) motion = F2{0, 0};
    L2 = F3{clampf(L2.x, 0.0f, 10.0f), clampf(L2.y, 0.0f, 10.0f), clampf(L2.z, 0.0f, 10.0f)};
    L2 = L2 / r.albedo;
    return PixelOut{L2, mask, outputNormal, r.albedo, outputDepth, motion};
}

}  // namespace orc

using namespace orc;

extern "C" {

struct orc_pt_args {
    const void* triangles; const void* nodes; const void* tlas; u32 numBatches; int stackSize, maxLoop;
    const void* materials; const int* materialsIdx; int numMaterials, numMaterialsIdx;
    const void* camera;            // 176-byte Camera blob
    const void* historyCamera;     // 176-byte Camera blob of the previous frame (HistoryCamera::Setup is applied here)
    const float* sky; const float* skyCdf; const float* sun; const float* sunCdf;
    float sunDir[3]; float sunAngleCosThetaMax;
    const uint8_t* blueNoise;      // sobol[65536] | scramble[131072] | rank[131072]
    const uint16_t* tex[2][11]; int texW[2][11], texH[2][11]; int texMaxLod[2];
    int frameNum, width, height;
    // outputs, float32, before the half stores of src/pathtrace.cuh:123-127
    float* color; uint16_t* mask; float* normal; float* albedo; float* depth; float* motion;
    unsigned long long* counters;  // [0] rays, [1] node visits, [2] leaf tests (may be NULL)
    int threads;
    int rowBegin, rowEnd;          // rows [rowBegin,rowEnd) are traced (bounded CPU-baseline samples)
};

int orc_pathtrace(const orc_pt_args* a) {
    PtScene s;
    s.bvh = Scene{(const Triangle*)a->triangles, (const BVHNode*)a->nodes, (const BVHNode*)a->tlas, a->numBatches, a->stackSize, a->maxLoop};
    s.materials = (const SurfaceMaterial*)a->materials; s.materialsIdx = a->materialsIdx;
    s.numMaterials = a->numMaterials; s.numMaterialsIdx = a->numMaterialsIdx;
    std::memcpy(&s.camera, a->camera, sizeof(Camera));
    Camera hc; std::memcpy(&hc, a->historyCamera, sizeof(Camera));
    s.historyCamera = history_setup(hc);
    s.env = Env{a->sky, a->skyCdf, a->sun, a->sunCdf, F3{a->sunDir[0], a->sunDir[1], a->sunDir[2]}, a->sunAngleCosThetaMax};
    s.bn = BlueNoise{a->blueNoise, a->blueNoise + 65536, a->blueNoise + 65536 + 131072};
    for (int t = 0; t < 2; ++t) {
        for (int l = 0; l < 11; ++l) { s.tex[t].mip[l] = a->tex[t][l]; s.tex[t].w[l] = a->texW[t][l]; s.tex[t].h[l] = a->texH[t][l]; }
        s.tex[t].maxLod = a->texMaxLod[t];
    }
    s.frameNum = a->frameNum;
    const int W = a->width, H = a->height;
    int threads = a->threads < 1 ? 1 : a->threads;
    std::vector<unsigned long long> cnt((size_t)threads * 3, 0);
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
        pool.emplace_back([&, t]() {
            for (int y = a->rowBegin + t; y < a->rowEnd; y += threads)
                for (int x = 0; x < W; ++x) {
                    PixelOut o = trace_pixel(s, x, y, &cnt[(size_t)t * 3]);
                    size_t i = (size_t)y * W + x;
                    a->color[3 * i] = o.color.x; a->color[3 * i + 1] = o.color.y; a->color[3 * i + 2] = o.color.z;
                    a->mask[i] = o.mask;
                    a->normal[3 * i] = o.normal.x; a->normal[3 * i + 1] = o.normal.y; a->normal[3 * i + 2] = o.normal.z;
                    a->albedo[3 * i] = o.albedo.x; a->albedo[3 * i + 1] = o.albedo.y; a->albedo[3 * i + 2] = o.albedo.z;
                    a->depth[i] = o.depth;
                    a->motion[2 * i] = o.motion.x; a->motion[2 * i + 1] = o.motion.y;
                }
        });
    for (auto& th : pool) th.join();
    if (a->counters) for (int t = 0; t < threads; ++t) for (int k = 0; k < 3; ++k) a->counters[k] += cnt[(size_t)t * 3 + k];
    (void)H;
    return 0;
}

}  // extern "C"
