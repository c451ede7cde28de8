// Device functions shared by the two path-trace drivers (megakernel: pathtrace.cu, wavefront: pathtrace_wf.cu).
// Semantics restated from src/pathtrace.cuh:11-128, src/raygen.cuh, src/traverse.cuh (RaySceneIntersect, UpdateMaterial),
// src/surfaceInteraction.cuh (Glossy/DiffuseSurfaceInteraction) and src/light.cuh:275-305.
#pragma once
#include "shading_bsdf.cuh"
#include "traverse.cuh"
#include "context.h"
#include <cuda_fp16.h>

namespace rtrt {

constexpr int kPtBlockX = 8, kPtBlockY = 8;       // src/kernel.cu:336: 8x8 pixel tiles (a warp covers 8x4 pixels)
constexpr int kPtThreads = kPtBlockX * kPtBlockY;

struct PathTraceParams {
    SceneView sc;
    const SurfaceMaterial* __restrict__ materials; const int* __restrict__ materialsIdx; int numMaterials, numMaterialsIdx;
    Camera camera; HistoryCamera historyCamera;
    EnvMaps env; BlueNoiseTables bn; MipTex tex[2];
    int frameNum, width, height, rowBegin, rowEnd;
    int stripH, stripShards, stripShard, localRows;        // row strips (stripShards > 1) or the contiguous band [rowBegin, rowEnd)
    int chunks, chunk, chunkRows;                          // this launch covers the local 8-row strips s with s % chunks == chunk
    ushort4* color; ushort4* normal; ushort4* albedo; __half* depth; __half2* motion;
    unsigned long long* rayCounter;
    int anyHitRays;                                        // option any_hit_rays (wavefront path): see kAnyHitBit in pathtrace_wf.cu
};

// row jj of this launch's chunk (0 .. chunkRows-1) -> local row of the context (0 .. localRows-1)
VD int chunk_row(const PathTraceParams& P, int jj) {
    if (P.chunks <= 1) return jj;
    return (((jj >> 3) * P.chunks + P.chunk) << 3) + (jj & 7);
}

// local row j (0 .. localRows-1) of this context -> frame row
VD int frame_row(const PathTraceParams& P, int j) {
    if (P.stripShards <= 1) return P.rowBegin + j;
    const int q = j / P.stripH;
    return (q * P.stripShards + P.stripShard) * P.stripH + (j - q * P.stripH);
}

struct RayState {            // live subset of src/kernel.cuh:233-284
    Float3 orig, dir, beta0, beta1, pos, normal, fakeNormal, albedo, centerRaydir;
    int matId, matType, lightIdx, objectIdx;
    float offset, normalDotRayDir, depth, rayConeWidth, rayConeSpread;
    bool isRayIntoSurface, hitLight, hit, isDiffuse, isHitProcessed, isOccluded, isShadowRay;
};

// src/raygen.cuh:45-63
VD float ray_cone_width(const Camera& cam, int ix, int iy) {
    const Float2 pc = (Float2{(float)ix, (float)iy} + 0.5f) - Float2{cam.resolution.x, cam.resolution.y} / 2.0f;
    const Float2 po{copysignf(0.5f, pc.x), copysignf(0.5f, pc.y)};
    const Float2 uvNear = (pc - po) * cam.inversedResolution * 2.0f;
    const Float2 uvFar = (pc + po) * cam.inversedResolution * 2.0f;
    const Float2 hf{tanf(cam.fov.x / 2.0f), tanf(cam.fov.y / 2.0f)};
    const Float2 n = uvNear * hf, f = uvFar * hf;
    return atanf(sqrtf(f.x * f.x + f.y * f.y)) - atanf(sqrtf(n.x * n.x + n.y * n.y));
}

// src/raygen.cuh:7-38
VD void generate_ray(Float3& orig, Float3& dir, Float3& centerDir, Float2& sampleUv, const Camera& cam, int ix, int iy, Float2 rp, Float2 ra) {
    Float2 uv = (Float2{(float)ix, (float)iy} + rp) * cam.inversedResolution;
    Float2 uvc = (Float2{(float)ix, (float)iy} + 0.5f) * cam.inversedResolution;
    sampleUv = uv;
    uv = uv * -2.0f + 1.0f;
    uvc = uvc * -2.0f + 1.0f;
    const Float3 p = cam.adjustedFront + cam.adjustedLeft * uv.x + cam.adjustedUp * uv.y;
    const Float3 pc = cam.adjustedFront + cam.adjustedLeft * uvc.x + cam.adjustedUp * uvc.y;
    const Float2 disk = concentric_sample_disk(ra);
    const Float3 pa = disk.x * cam.apertureLeft + disk.y * cam.apertureUp;
    orig = cam.pos + pa;
    dir = normalize(p - pa);
    centerDir = normalize(pc);
}

// src/traverse.cuh:9-56
VD void update_material(const PathTraceParams& P, RayState& r) {
    if (!r.hit) { r.matType = MAT_SKY; r.matId = 99999; }
    else {
        r.matId = (r.objectIdx >= 0 && r.objectIdx < P.numMaterialsIdx) ? __ldg(P.materialsIdx + r.objectIdx) : 6;   // SAFE_LOAD default 6
        r.matType = (r.matId >= 0 && r.matId < P.numMaterials) ? (int)P.materials[r.matId].type : (int)PERFECT_REFLECTION;
    }
    if (r.isShadowRay) {
        if ((r.matType == EMISSIVE && r.lightIdx == r.objectIdx) || (r.matType == MAT_SKY && r.lightIdx == kEnvLightId)) r.hitLight = true;
        else r.isOccluded = true;
    } else {
        r.hitLight = r.matType == EMISSIVE || r.matType == MAT_SKY;
    }
    r.isDiffuse = (r.matType == LAMBERTIAN_DIFFUSE) || (r.matType == MICROFACET_REFLECTION);
}

// src/traverse.cuh:64-225 split in two so that the wavefront driver can run the traversal in a separate kernel:
// needs_ray = the guard at the top of RaySceneIntersect; apply_hit = everything after the TraverseBvh call
// (hit record src/traverse.h:161-174, normal flipping, hit flag, ray-cone growth, UpdateMaterial).
VD bool needs_ray(const RayState& r) { return !(r.hitLight || !r.isHitProcessed || r.isOccluded); }

VD void apply_hit(const PathTraceParams& P, RayState& r, float ht, int htri, float hua, float hva) {
    r.objectIdx = htri;
    r.offset = 1e-7f;
    r.normal = f3(0.0f);
    r.pos = f3(kRayMax);
    if (htri >= 0) {
        const float4* row = P.sc.triRows + (size_t)htri * 8;
        const float4 q1 = __ldg(row), q2 = __ldg(row + 1), q3 = __ldg(row + 2);
        const float4 m1 = __ldg(row + 4), m2 = __ldg(row + 5), m3 = __ldg(row + 6);
        const Float3 v1{q1.x, q1.y, q1.z}, v2{q2.x, q2.y, q2.z}, v3{q3.x, q3.y, q3.z};
        const Float3 n = normalize(cross(v2 - v1, v3 - v1));
        const float errorT = (16 * 1.1920928955078125e-07f) / (1.0f - 16 * 1.1920928955078125e-07f) * fabsf(ht);   // ErrGamma(16)*|t|
        // GetRayPlaneIntersectPoint (src/geometry.cuh:252-262)
        const Float3 p = r.orig + r.dir * ht;
        const Float3 pc = p - (dot(n, p) + (-dot(n, v1))) * n;
        const float errorP = getmax3(abs3(pc)) * ((6 * 1.1920928955078125e-07f) / (1.0f - 6 * 1.1920928955078125e-07f));
        r.normal = n; r.pos = pc; r.offset = errorT + errorP;
        const Float3 n1 = normalize(Float3{m1.x, m1.y, m1.z}), n2 = normalize(Float3{m2.x, m2.y, m2.z}), n3 = normalize(Float3{m3.x, m3.y, m3.z});
        r.fakeNormal = normalize(n3 * (1.0f - hua - hva) + n1 * hua + n2 * hva);
    }
    r.normalDotRayDir = dot(r.normal, r.dir);
    r.isRayIntoSurface = r.normalDotRayDir < 0;
    if (!r.isRayIntoSurface) { r.normal = -r.normal; r.normalDotRayDir = -r.normalDotRayDir; }
    if (dot(r.fakeNormal, r.normal) < 0) r.fakeNormal = -r.fakeNormal;
    r.hit = (ht < kRayMax);
    r.depth = ht;
    if (!r.hit) { r.normal = f3(0, -1, 0); r.fakeNormal = f3(0, -1, 0); }
    else r.rayConeWidth += r.rayConeSpread * ht;
    update_material(P, r);
}

template <bool COUNT>
VD void ray_scene_intersect(const PathTraceParams& P, const AABB& sceneBox, RayState& r, uint32_t* stackLink, float* stackT, uint32_t& rays,
                            uint32_t& nodeVisits, uint32_t& leafTests) {
    if (!needs_ray(r)) return;
    r.isHitProcessed = false;
    ++rays;
    const RayPrep rp = prepare_ray(r.orig, r.dir, sceneBox);
    const TraverseResult h = traverse_bvh<16, COUNT>(P.sc, rp, 0, stackLink, stackT, kPtThreads);
    if (COUNT) { nodeVisits += h.nodeVisits; leafTests += h.leafTests; }
    apply_hit(P, r, h.t, h.tri, h.ua, h.va);
}

// src/surfaceInteraction.cuh:11-34
VD void glossy_interaction(RayState& r, float rnd) {
    if (r.hitLight || r.isDiffuse || r.isOccluded) return;
    r.isHitProcessed = true;
    if (r.matType == PERFECT_REFLECTION) {
        r.dir = normalize(r.dir - r.normal * dot(r.dir, r.normal) * 2.0f);
        r.orig = r.pos + r.offset * r.normal;
    } else if (r.matType == PERFECT_FRESNEL_REFLECTION_REFRACTION) {
        Float3 next; float off = r.offset;
        perfect_reflection_refraction(1.0f, 1.33f, r.isRayIntoSurface, r.normal, r.normalDotRayDir, rnd, r.dir, next, off);
        r.dir = next;
        r.orig = r.pos + off * r.normal;
    }
}

// Texture gamma (src/surfaceInteraction.cuh:118-120: pow3f(albedo, 2.2f)).  powf is ~150 SASS instructions per call and runs nine
// times per shaded hit; v^a = 2^(a * log2 v) on the SFU (MUFU.LG2 / MUFU.EX2) is three.  The operand is a convex combination of
// unorm16 texels (>= 0; log2 0 = -inf gives 0 like powf), the relative error is <= |a * log2 v| * 2^-22 + 2^-22 < 1e-6 for the
// texel range, against the stated 1e-3 rel-RMSE radiance tolerance; the value feeds throughput only, never a branch.
VD float pow_sfu(float v, float a) {
    float l, r;
    asm("lg2.approx.f32 %0, %1;" : "=f"(l) : "f"(v));
    asm("ex2.approx.f32 %0, %1;" : "=f"(r) : "f"(a * l));
    return r;
}
VD Float3 pow3f(Float3 v, float a) { return Float3{pow_sfu(v.x, a), pow_sfu(v.y, a), pow_sfu(v.z, a)}; }

// SampleBicubicSmoothStepLod (src/sampler.cuh:565-584) with the two mip levels as a rolled loop (code size)
VD Float4 sample_tex_lod_rolled(const MipTex* texs, int t, Float2 uv, float lod) {
    const float low = floorf(lod);
    const float f = lod - low;
    int lo = __float2int_rz(low);                 // cvt.rzi.s32.f32: saturating, NaN -> 0
    int hi = (lo == 0x7fffffff) ? lo : lo + 1;
    const int maxLod = texs[t].maxLod;
    lo = clampi(lo, 0, maxLod); hi = clampi(hi, 0, maxLod);
    Float4 acc{0, 0, 0, 0};
#pragma unroll 1
    for (int k = 0; k < 2; ++k) {
        const int lv = k ? hi : lo;
        const int w = texs[t].w[lv], h = texs[t].h[lv];
        const ushort4* __restrict__ img = texs[t].mip[lv];
        const BicubicTaps tp = bicubic_taps(uv, w, h);
        int xs[2], ys[2];
        repeat_pair(tp.x0, w, xs[0], xs[1]);
        repeat_pair(tp.y0, h, ys[0], ys[1]);
        Float4 out{0, 0, 0, 0};
#pragma unroll
        for (int i = 0; i < 4; ++i) out = out + load_ushort4(img, w, h, xs[i & 1], ys[i >> 1]) * tp.w[i];
        out = out / tp.sum;
        acc = k ? (acc + out * f) : (out * (1.0f - f));       // a * (1 - f) + b * f
    }
    return acc;
}

// one triplanar projection of src/surfaceInteraction.cuh:96-155
VD void triplanar_axis(const PathTraceParams& P, Float2 uv, float lod, Float3 normal, Float3 w, Float3& albedo, Float3& nrm) {
    uv = uv * 0.5f;
    Float4 c{0, 0, 0, 0}, c2{0, 0, 0, 0};
#pragma unroll 1
    for (int t = 0; t < 2; ++t) {
        const Float4 v = sample_tex_lod_rolled(P.tex, t, uv, lod);
        if (t == 0) c = v; else c2 = v;
    }
    albedo = pow3f(xyz(c), 2.2f);
    const Float3 n = xyz(c2) - 0.5f;
    const Float3 u = cross(normal, w);
    const Float3 v = cross(normal, u);
    nrm = normalize(u * n.x + v * n.y + normal * n.z);
}

// src/surfaceInteraction.cuh:36-310
VD void diffuse_interaction(int bounce, const PathTraceParams& P, RayState& r, Float3& beta, const Float4& rnd, const Float4& rnd2) {
    if (r.hitLight || !r.isDiffuse || r.isOccluded) return;
    r.lightIdx = kDefaultLightId;
    r.isHitProcessed = true;
    const SurfaceMaterial mat = P.materials[r.matId];
    Float3 normal = r.fakeNormal;
    const Float3 surfaceNormal = r.normal;
    Float3 albedo;
    {
        const float tw = (float)P.tex[0].w[0], th = (float)P.tex[0].h[0];
        const float lod = log2f(r.rayConeWidth * 0.5f * sqrtf(tw * tw + th * th));
        const float wx = surfaceNormal.x * surfaceNormal.x, wy = surfaceNormal.y * surfaceNormal.y, wz = surfaceNormal.z * surfaceNormal.z;
        const Float3 shadingN = normal;
        Float3 aSum = f3(0.0f), nSum = f3(0.0f);
#pragma unroll 1
        for (int ax = 0; ax < 3; ++ax) {                       // X, Y, Z projections; sums keep the reference's left-to-right order
            const Float2 uv = ax == 0 ? Float2{r.pos.y, r.pos.z} : (ax == 1 ? Float2{r.pos.x, r.pos.z} : Float2{r.pos.x, r.pos.y});
            const Float3 w = ax == 1 ? (fabsf(shadingN.x) > 0.999f ? f3(0, 0, 1) : f3(1, 0, 0))
                                     : (fabsf(shadingN.y) > 0.999f ? (ax == 0 ? f3(0, 0, 1) : f3(1, 0, 0)) : f3(0, 1, 0));
            const float wt = ax == 0 ? wx : (ax == 1 ? wy : wz);
            Float3 a, n;
            triplanar_axis(P, uv, lod, shadingN, w, a, n);
            if (ax == 0) { aSum = a * wt; nSum = n * wt; } else { aSum = aSum + a * wt; nSum = nSum + n * wt; }
        }
        albedo = aSum;
        normal = normalize(nSum);
        r.fakeNormal = normal;
    }
    if (bounce == 0) r.albedo = albedo * (1.0f + fabsf(dot(normal, r.centerRaydir)));
    const Float3 rayDir = r.dir;
    Float3 lightDir; float lightPdf = 1.0f; int lightIdx;
    sample_light(P.env, lightDir, lightPdf, lightIdx, rnd2);
    Float3 surfDir = f3(0.0f), sBsdf = f3(0.0f), lBsdf = f3(0.0f);
    float sPdf = 0.0f, lPdf = 0.0f;
    const Float2 r2{rnd.x, rnd.y}, r22{rnd.z, rnd.w};
    if (r.matType == LAMBERTIAN_DIFFUSE) {
        surfDir = lambertian_sample(r2, normal);
        sBsdf = lambertian_bsdf(albedo); sPdf = lambertian_pdf(surfDir, normal);
        lBsdf = lambertian_bsdf(albedo);
    } else if (r.matType == MICROFACET_REFLECTION) {
        microfacet_reflection_sample(r2, r22, rayDir, surfDir, normal, surfaceNormal, sBsdf, sPdf, mat.F0, albedo, mat.alpha);
        microfacet_reflection(lBsdf, lPdf, normal, -rayDir, lightDir, mat.F0, albedo, mat.alpha);
    }
    if (isnan3(lBsdf)) lBsdf = f3(0.0f);                    // NAN_DETECTER src/surfaceInteraction.cuh:228-231
    if (isnan3(sBsdf)) sBsdf = f3(0.0f);
    if (sPdf != sPdf) sPdf = 0.0f;
    if (lightPdf != lightPdf) lightPdf = 0.0f;
    const float powerHeuristicSurface = (sPdf * sPdf) / (sPdf * sPdf + lightPdf * lightPdf);
    const float minFinalPdf = 1e-5f;
    if (rnd.w < powerHeuristicSurface) {
        if (dot(r.normal, surfDir) < 0) { r.isOccluded = true; return; }
        const float cosWi = sel_max(RTRT_SAFE_COSINE_EPSI, dot(surfDir, normal));
        beta = sBsdf * cosWi / sel_max(sPdf, minFinalPdf);
        r.dir = surfDir;
    } else {
        if (dot(r.normal, lightDir) < 0) { r.isOccluded = true; return; }
        const float cosWi = sel_max(RTRT_SAFE_COSINE_EPSI, dot(lightDir, normal));
        beta = lBsdf * cosWi / sel_max(lightPdf, minFinalPdf);
        r.dir = lightDir;
        r.lightIdx = lightIdx;
        r.isShadowRay = true;
    }
    if (isnan3(beta)) beta = f3(0.0f);
    r.orig = r.pos + r.offset * r.normal;
}

VD unsigned short f2h_bits(float f) { return __half_as_ushort(__float2half_rn(f)); }

}  // namespace rtrt
