// Frame resources of the hot path: G-buffers, blue-noise tables, materials, texture atlas, and the sky/sun
// environment maps + luminance CDFs the path tracer samples (the input-prep block of RayTracer::draw,
// src/kernel.cu:285-308: Sky -> Scan -> SkySun -> Scan, re-run only when the sky parameters change).
//   src/sky.cuh:91-319     Hosek-Wilkie sky radiance (10 spectral bands), solar disc radiance, Sky / SkySun kernels
//   src/kernel.cu:121-124  sun direction from time of day
//   src/scan.cuh           inclusive scan of the luminance PDFs (here: one CTA, shuffle scan)
//   src/init.cu:215-265    default materials;  src/init.cu:441-584 buffers;  src/mipgen.cu:121-139 mip rule
#include "context.h"
#include "rtrt_vec.cuh"
#include <dlfcn.h>
#include <vector>
#include <cstring>
#include <cmath>

namespace rtrt {

// ------------------------------------------------------------------------------------------------
// sky / sun maps
// ------------------------------------------------------------------------------------------------
struct SkyConsts {
    float cfg[90]; float rad[10];          // per-band fitted coefficients (UpdateSkyState, src/sky.cuh:126-148)
    Float3 sunDir; float skyScalar, sunScalar, sunAngle, cosThetaMax;
};

__constant__ float kCieX[10] = {2.372527e-02f, 1.955480e+00f, 1.074553e+01f, 5.056697e+00f, 4.698190e+00f, 2.391135e+01f, 3.798705e+01f, 1.929414e+01f, 2.970610e+00f, 2.092986e-01f};
__constant__ float kCieY[10] = {6.813859e-04f, 6.771017e-02f, 1.171193e+00f, 6.997765e+00f, 2.666710e+01f, 3.758372e+01f, 2.503930e+01f, 8.150395e+00f, 1.098635e+00f, 7.563256e-02f};
__constant__ float kCieZ[10] = {1.119121e-01f, 9.441195e+00f, 5.597921e+01f, 3.589996e+01f, 5.070894e+00f, 3.523189e-01f, 3.422707e-02f, 2.539118e-03f, 7.836666e-06f, 0.000000e+00f};

VD Float3 xyz_to_srgb(Float3 c) {         // src/color.h:19-30
    return Float3{dot(Float3{3.2404542f, -1.5371385f, -0.4985314f}, c), dot(Float3{-0.9692660f, 1.8760108f, 0.0415560f}, c),
                  dot(Float3{0.0556434f, -0.2040259f, 1.0572252f}, c)};
}

// src/sky.cuh:166-197
VD Float3 sky_radiance(Float3 raydir, const SkyConsts& k) {
    const float theta = acosf(raydir.y);
    const float gamma = acosf(clampf(dot(raydir, k.sunDir), -1.0f, 1.0f));
    Float3 xyzc = f3(0.0f);
#pragma unroll 1
    for (int ch = 0; ch < 10; ++ch) {
        const float* c = k.cfg + ch * 9;
        const float expM = expf(c[4] * gamma);
        const float rayM = cosf(gamma) * cosf(gamma);
        const float mieM = (1.0f + cosf(gamma) * cosf(gamma)) / powf((1.0f + c[8] * c[8] - 2.0f * c[8] * cosf(gamma)), 1.5f);
        const float zenith = sqrtf(cosf(theta));
        const float ri = (1.0f + c[0] * expf(c[1] / (cosf(theta) + 0.01f))) * (c[2] + c[3] * expM + c[5] * rayM + c[6] * mieM + c[7] * zenith);
        const float radiance = ri * k.rad[ch];
        xyzc = xyzc + radiance * Float3{kCieX[ch], kCieY[ch], kCieZ[ch]};
    }
    return xyz_to_srgb(xyzc);
}

// src/sky.cuh:199-277 ; hosek = skyDataSets[540] | skyDataSetsRad[60] | solarDatasets[1800] | limbDarkening[60]
VD Float3 sun_radiance(Float3 raydir, const SkyConsts& k, const float* __restrict__ hosek) {
    const float* solar = hosek + 600; const float* limb = hosek + 2400;
    const float gamma = acosf(clampf(dot(raydir, k.sunDir), -1.0f, 1.0f));
    const float elevation = (RTRT_M_PI / 2.0f) - acosf(k.sunDir.y);
    const float solarRadius = k.sunAngle * RTRT_M_PI / 180.0f / 2.0f;
    const float scale = 1.0f / ((k.sunAngle / 0.51f) * (k.sunAngle / 0.51f));
    const float srs = sinf(solarRadius);
    const float ar2 = 1.0f / (srs * srs);
    const float sg = sinf(gamma);
    float sc2 = 1.0f - ar2 * sg * sg;
    if (sc2 < 0.0f) sc2 = 0.0f;
    const float sampleCosine = sqrtf(sc2);
    if (sampleCosine == 0.0f) return f3(0.0f);
    Float3 xyzc = f3(0.0f);
#pragma unroll 1
    for (int ch = 0; ch < 10; ++ch) {
        const int pieces = 45, order = 4;
        int pos = (int)(powf((float)(2.0 * elevation / RTRT_M_PI), (float)(1.0 / 3.0)) * pieces);
        if (pos > 44) pos = 44;
        const float break_x = (float)(powf(((float)pos / (float)pieces), 3.0f) * ((double)RTRT_M_PI * 0.5));
        const float* coefs = solar + ch * 180 + (order * (pos + 1) - 1);
        float res = 0.0f;
        const float x = elevation - break_x;
        float x_exp = 1.0f;
        for (int i = 0; i < order; ++i) { res += x_exp * __ldg(coefs - i); x_exp *= x; }
        const float* ld = limb + ch * 6;
        const float dark = __ldg(ld) + __ldg(ld + 1) * sampleCosine + __ldg(ld + 2) * powf(sampleCosine, 2.0f) + __ldg(ld + 3) * powf(sampleCosine, 3.0f) +
                           __ldg(ld + 4) * powf(sampleCosine, 4.0f) + __ldg(ld + 5) * powf(sampleCosine, 5.0f);
        const float direct = res * (dark * scale);
        xyzc = xyzc + direct * Float3{kCieX[ch], kCieY[ch], kCieZ[ch]};
    }
    return xyz_to_srgb(xyzc);
}

// Sky (src/sky.cuh:280-299) and SkySun (:301-319) in one launch: blocks [0, skyBlocks) fill the 512x256 sky map,
// the rest the 32x32 sun-cone map.  Each writes the float4 map and the luminance PDF.
__global__ void k_sky_maps(const SkyConsts k, const float* __restrict__ hosek, float4* sky, float* skyPdf, float4* sun, float* sunPdf) {
    const int gid = blockIdx.x * blockDim.x + threadIdx.x;
    if (gid < kSkySize) {
        const int x = gid % kSkyW, y = gid / kSkyW;
        const float u = ((float)x + 0.5f) / kSkyW, v = ((float)y + 0.5f) / kSkyH;
        Float3 c = sky_radiance(equal_area_map(u, v), k) * k.skyScalar;
        c = max3f(c, f3(0.0f));
        sky[gid] = make_float4(c.x, c.y, c.z, 0.0f);
        skyPdf[gid] = dot(c, Float3{0.3f, 0.6f, 0.1f});
    } else if (gid < kSkySize + kSunSize) {
        const int i = gid - kSkySize;
        const int x = i % kSunW, y = i / kSunW;
        const float u = ((float)x + 0.5f) / kSunW, v = ((float)y + 0.5f) / kSunH;
        const Float3 dir = equal_area_map_cone(k.sunDir, u, v, k.cosThetaMax);
        Float3 c = sun_radiance(dir, k, hosek) * k.sunScalar;
        c = max3f(c, f3(0.0f));
        sun[i] = make_float4(c.x, c.y, c.z, 0.0f);
        sunPdf[i] = dot(c, Float3{0.3f, 0.6f, 0.1f});
    }
}

// Inclusive scan (postfix = 1) with EXACTLY the summation order of the reference's CpuScan (src/scan.cuh:235-251):
// out[i] = ((in[0] + in[1]) + ...) + in[i].  Float addition is not associative and the CDF must be monotone for
// the binary search in SampleLight, so the order is kept sequential; one warp per array streams 32 values per
// coalesced load and every lane replays the same running sum through shuffles.  Runs only when the sky changes.
__global__ void __launch_bounds__(32) k_inclusive_scan(const float* __restrict__ in0, float* out0, int n0, const float* __restrict__ in1, float* out1, int n1) {
    const float* in = blockIdx.x == 0 ? in0 : in1; float* out = blockIdx.x == 0 ? out0 : out1; const int n = blockIdx.x == 0 ? n0 : n1;
    const int lane = threadIdx.x;
    float acc = 0.0f;
    for (int base = 0; base < n; base += 32) {
        const float v = (base + lane < n) ? in[base + lane] : 0.0f;
        float mine = 0.0f;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            acc += __shfl_sync(0xffffffffu, v, j);
            if (j == lane) mine = acc;
        }
        if (base + lane < n) out[base + lane] = mine;
    }
}

// host side of the regeneration: sun direction (src/kernel.cu:121-124) and UpdateSkyState (src/sky.cuh:126-148)
static Float3 h_norm(Float3 v) { const float n = sqrtf(v.x * v.x + v.y * v.y + v.z * v.z); return Float3{v.x / n, v.y / n, v.z / n}; }
static Float3 h_cross(Float3 a, Float3 b) {
    auto dop = [](float a, float b, float c, float d) { const float cd = c * d; const float err = fmaf(-c, d, cd); const float r = fmaf(a, b, -cd); return r + err; };
    return Float3{dop(a.y, b.z, a.z, b.y), dop(a.z, b.x, a.x, b.z), dop(a.x, b.y, a.y, b.x)};
}
static float h_dot(Float3 a, Float3 b) { return (float)((double)a.x * b.x + (double)a.y * b.y + (double)a.z * b.z); }
static Float3 h_add(Float3 a, Float3 b) { return Float3{a.x + b.x, a.y + b.y, a.z + b.z}; }
static Float3 h_scale(Float3 a, float s) { return Float3{a.x * s, a.y * s, a.z * s}; }

static Float3 compute_sun_dir(float timeOfDay, float sunAxisAngleDeg) {
    const float d2r = 0.01745329251f;
    const Float3 axis = h_norm(Float3{0.0f, cosf(sunAxisAngleDeg * d2r), sinf(sunAxisAngleDeg * d2r)});
    const float angle = fmodf(timeOfDay * RTRT_M_PI, RTRT_TWO_PI);
    const Float3 v = h_cross(Float3{0, 1, 0}, axis);
    const Float3 qv = h_scale(h_norm(axis), sinf(angle / 2)); const float qw = cosf(angle / 2);   // Quat::axisAngle
    const Float3 pv = h_add(h_scale(v, qw), h_cross(qv, v)); const float pw = -h_dot(qv, v);         // q * (v,0)
    const Float3 cv{-qv.x, -qv.y, -qv.z};                                                             // * conj(q)
    const Float3 rv = h_add(h_add(h_scale(cv, pw), h_scale(pv, qw)), h_cross(pv, cv));
    return h_norm(rv);
}

static float fit5(const float* m, float e, int i, int stride) {
    return (powf(1.0f - e, 5.0f) * m[i] + 5.0f * powf(1.0f - e, 4.0f) * e * m[i + stride] + 10.0f * powf(1.0f - e, 3.0f) * powf(e, 2.0f) * m[i + 2 * stride] +
            10.0f * powf(1.0f - e, 2.0f) * powf(e, 3.0f) * m[i + 3 * stride] + 5.0f * (1.0f - e) * powf(e, 4.0f) * m[i + 4 * stride] + powf(e, 5.0f) * m[i + 5 * stride]);
}

static int regenerate_sky(Context& c, const std::vector<float>& hosek) {
    rtrt_sky_params p = c.skyParams;
    p.sunScalar = fmaxf(p.sunScalar, 0.00001f); p.skyScalar = fmaxf(p.skyScalar, 0.00001f); p.sunAngle = fmaxf(p.sunAngle, 0.51f);   // src/kernel.cu:291-293
    c.sunDir = compute_sun_dir(p.timeOfDay, p.sunAxisAngle);
    SkyConsts k{};
    const float elevation = acosf(c.sunDir.y);
    const float se = powf(elevation / (RTRT_M_PI / 2.0f), (1.0f / 3.0f));
    for (int ch = 0; ch < 10; ++ch) {
        for (int i = 0; i < 9; ++i) k.cfg[ch * 9 + i] = fit5(hosek.data() + ch * 54, se, i, 9);
        k.rad[ch] = fit5(hosek.data() + 540 + ch * 6, se, 0, 1);
    }
    k.sunDir = c.sunDir; k.skyScalar = p.skyScalar; k.sunScalar = p.sunScalar; k.sunAngle = p.sunAngle;
    k.cosThetaMax = cosf(p.sunAngle * RTRT_M_PI / 180.0f / 2.0f);
    c.sunAngleCosThetaMax = k.cosThetaMax;                                                         // src/kernel.cu:303-305
    const int total = kSkySize + kSunSize;
    k_sky_maps<<<(total + 127) / 128, 128, 0, c.stream>>>(k, c.dHosek, reinterpret_cast<float4*>(c.dSky), c.dSkyPdf,
                                                         reinterpret_cast<float4*>(c.dSun), c.dSunPdf);
    k_inclusive_scan<<<2, 32, 0, c.stream>>>(c.dSkyPdf, c.dSkyCdf, kSkySize, c.dSunPdf, c.dSunCdf, kSunSize);
    c.launches += 2;
    RTRT_CUDA(c, cudaGetLastError());
    c.skyDirty = false;
    return RTRT_OK;
}

}  // namespace rtrt

// ------------------------------------------------------------------------------------------------
// resources
// ------------------------------------------------------------------------------------------------
namespace rtrt {

static std::string default_data_dir() {
    if (const char* e = getenv("RTRT_DATA_DIR")) return e;
    Dl_info info;
    if (dladdr((void*)&default_data_dir, &info) && info.dli_fname) {
        std::string p = info.dli_fname;
        const size_t s = p.find_last_of('/');
        return (s == std::string::npos ? std::string(".") : p.substr(0, s)) + "/data";
    }
    return "data";
}

static bool read_file(const std::string& path, void* dst, size_t bytes) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return false;
    const size_t n = fread(dst, 1, bytes, f);
    fclose(f);
    return n == bytes;
}

template <typename T>
static void dfree2(T*& p) { if (p) cudaFree(p); p = nullptr; }

static void free_gbuffers(Context& c) {
    void** bufs[] = {&c.dColor, &c.dNormal, &c.dAlbedo, &c.dDepth, &c.dMotion, &c.dAccum, &c.dHistoryColor, &c.dHistoryDepth,
                     &c.dHistoryAlbedo, &c.dNoise, &c.dNoise16, &c.dColorTmp, &c.dGuide};
    for (void** b : bufs) { if (*b) cudaFree(*b); *b = nullptr; }
    if (c.dAccumRadiance) { cudaFree(c.dAccumRadiance); c.dAccumRadiance = nullptr; c.accumRadiancePx = 0; }
    c.gbufW = c.gbufH = c.gbufCapW = c.gbufCapH = 0;
    c.accumW = c.accumH = c.histW = c.histH = 0;
}

void free_frame_resources(Context& c) {
    free_gbuffers(c);
    dfree2(c.dHosek); dfree2(c.dBlueNoise);
    dfree2(c.dSky); dfree2(c.dSkyPdf); dfree2(c.dSkyCdf); dfree2(c.dSun); dfree2(c.dSunPdf); dfree2(c.dSunCdf);
    dfree2(c.dMaterials); dfree2(c.dMaterialsIdx);
    for (int t = 0; t < 2; ++t) { if (c.dTexMem[t]) cudaFree(c.dTexMem[t]); c.dTexMem[t] = nullptr; }
    dfree2(c.dRayCounter);
    if (c.dWf) { cudaFree(c.dWf); c.dWf = nullptr; c.wfBytes = 0; }
    dfree2(c.dOutRgba); dfree2(c.dColor64); dfree2(c.dExposure);
}

// Texture slot upload + mip chain (2x2 box in float, truncating store: MipmapGen src/mipgen.cu:121-139; init-time work)
int set_texture(Context& c, int slot, int w, int h, const uint16_t* rgba) {
    if (slot < 0 || slot > 1 || w < 1 || h < 1 || !rgba) return c.fail(RTRT_ERR_ARG, "rtrt_set_texture: bad argument");
    std::vector<std::vector<uint16_t>> mips;
    std::vector<int> ws, hs;
    mips.emplace_back(rgba, rgba + (size_t)4 * w * h); ws.push_back(w); hs.push_back(h);
    while ((ws.back() != 1 || hs.back() != 1) && (int)mips.size() < kMaxMips) {
        const int pw = ws.back(), ph = hs.back(), nw = pw / 2 > 1 ? pw / 2 : 1, nh = ph / 2 > 1 ? ph / 2 : 1;
        const std::vector<uint16_t>& src = mips.back();
        std::vector<uint16_t> dst((size_t)4 * nw * nh);
        auto at = [&](int x, int y, int ch) { x = x < pw ? x : pw - 1; y = y < ph ? y : ph - 1; return (float)src[4 * ((size_t)y * pw + x) + ch]; };
        for (int y = 0; y < nh; ++y)
            for (int x = 0; x < nw; ++x)
                for (int ch = 0; ch < 4; ++ch) {
                    float v = (at(2 * x, 2 * y, ch) + at(2 * x + 1, 2 * y, ch) + at(2 * x, 2 * y + 1, ch) + at(2 * x + 1, 2 * y + 1, ch)) / 4.0f;
                    v = fminf(v, 65535.0f);
                    dst[4 * ((size_t)y * nw + x) + ch] = (uint16_t)v;
                }
        mips.push_back(std::move(dst)); ws.push_back(nw); hs.push_back(nh);
    }
    size_t total = 0;
    for (auto& m : mips) total += m.size();
    RTRT_CUDA(c, cudaSetDevice(c.device));
    RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    if (c.dTexMem[slot]) { cudaFree(c.dTexMem[slot]); c.dTexMem[slot] = nullptr; }
    RTRT_CUDA(c, cudaMalloc(&c.dTexMem[slot], total * sizeof(uint16_t)));
    MipTex& t = c.tex[slot];
    size_t off = 0;
    for (int l = 0; l < kMaxMips; ++l) {
        const int s = l < (int)mips.size() ? l : (int)mips.size() - 1;        // levels past the chain alias the last one
        if (l < (int)mips.size()) {
            RTRT_CUDA(c, cudaMemcpy((uint16_t*)c.dTexMem[slot] + off, mips[l].data(), mips[l].size() * sizeof(uint16_t), cudaMemcpyHostToDevice));
            t.mip[l] = reinterpret_cast<const ushort4*>((uint16_t*)c.dTexMem[slot] + off);
            off += mips[l].size();
        } else t.mip[l] = t.mip[s];
        t.w[l] = ws[s]; t.h[l] = hs[s];
    }
    t.maxLod = 10;                                                             // src/init.cu:577
    return RTRT_OK;
}

int set_materials(Context& c, const void* materials, int numMaterials, const int* materialsIdx, int numIdx) {
    RTRT_CUDA(c, cudaSetDevice(c.device));
    RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    dfree2(c.dMaterials); dfree2(c.dMaterialsIdx);
    std::vector<SurfaceMaterial> def;
    std::vector<int> idx;
    if (!materials) {
        // the reference scene: src/init.cu:215-251 (10 materials; default-constructed entries are the SurfaceMaterial() defaults)
        SurfaceMaterial d{Float3{0.8f, 0.8f, 0.8f}, PERFECT_REFLECTION, Float3{0.56f, 0.57f, 0.58f}, 0.05f};
        def.assign(10, d);
        def[0].type = EMISSIVE; def[0].albedo = Float3{0.1f, 0.2f, 0.9f};
        def[1].type = PERFECT_FRESNEL_REFLECTION_REFRACTION;
        def[2].type = EMISSIVE; def[2].albedo = Float3{0.9f, 0.2f, 0.1f};
        def[3].type = LAMBERTIAN_DIFFUSE; def[3].albedo = Float3{0.9f, 0.9f, 0.9f};
        def[4].type = MICROFACET_REFLECTION; def[4].albedo = Float3{0.9f, 0.9f, 0.9f};
        def[5].type = PERFECT_REFLECTION;
        def[6].type = LAMBERTIAN_DIFFUSE;
        def[7].type = LAMBERTIAN_DIFFUSE; def[7].albedo = Float3{0.9f, 0.2f, 0.1f};
        def[8].type = LAMBERTIAN_DIFFUSE; def[8].albedo = Float3{0.2f, 0.9f, 0.1f};
        def[9].type = LAMBERTIAN_DIFFUSE; def[9].albedo = Float3{0.1f, 0.2f, 0.9f};
        materials = def.data(); numMaterials = 10;
    }
    if (!materialsIdx) { idx.assign(c.numTris ? c.numTris : 1, 3); materialsIdx = idx.data(); numIdx = (int)idx.size(); }   // src/init.cu:262-265
    RTRT_CUDA(c, cudaMalloc((void**)&c.dMaterials, sizeof(SurfaceMaterial) * numMaterials));
    RTRT_CUDA(c, cudaMalloc((void**)&c.dMaterialsIdx, sizeof(int) * numIdx));
    RTRT_CUDA(c, cudaMemcpy(c.dMaterials, materials, sizeof(SurfaceMaterial) * numMaterials, cudaMemcpyHostToDevice));
    RTRT_CUDA(c, cudaMemcpy(c.dMaterialsIdx, materialsIdx, sizeof(int) * numIdx, cudaMemcpyHostToDevice));
    c.numMaterials = numMaterials; c.numMaterialsIdx = numIdx;
    return RTRT_OK;
}

int ensure_frame_resources(Context& c) {
    RTRT_CUDA(c, cudaSetDevice(c.device));
    if (c.width <= 0 || c.height <= 0) return c.fail(RTRT_ERR_STATE, "render size not set");
    if (c.dataDir.empty()) c.dataDir = default_data_dir();
    if (!c.dRayCounter) { RTRT_CUDA(c, cudaMalloc((void**)&c.dRayCounter, 64)); RTRT_CUDA(c, cudaMemsetAsync(c.dRayCounter, 0, 64, c.stream)); }
    if (!c.dBlueNoise) {
        std::vector<uint8_t> bn(65536 + 131072 + 131072);
        if (!read_file(c.dataDir + "/blue_noise_4spp.bin", bn.data(), bn.size())) return c.fail(RTRT_ERR_IO, "cannot read data/blue_noise_4spp.bin (rtrt_set_data_dir?)");
        RTRT_CUDA(c, cudaMalloc((void**)&c.dBlueNoise, bn.size()));
        RTRT_CUDA(c, cudaMemcpy(c.dBlueNoise, bn.data(), bn.size(), cudaMemcpyHostToDevice));
    }
    if (!c.dSky) {
        RTRT_CUDA(c, cudaMalloc((void**)&c.dSky, sizeof(float) * 4 * kSkySize)); RTRT_CUDA(c, cudaMalloc((void**)&c.dSkyPdf, sizeof(float) * kSkySize));
        RTRT_CUDA(c, cudaMalloc((void**)&c.dSkyCdf, sizeof(float) * kSkySize)); RTRT_CUDA(c, cudaMalloc((void**)&c.dSun, sizeof(float) * 4 * kSunSize));
        RTRT_CUDA(c, cudaMalloc((void**)&c.dSunPdf, sizeof(float) * kSunSize)); RTRT_CUDA(c, cudaMalloc((void**)&c.dSunCdf, sizeof(float) * kSunSize));
        c.skyDirty = true;
    }
    if (c.skyDirty) {
        std::vector<float> hosek(2460);
        if (!read_file(c.dataDir + "/hosek_sky.bin", hosek.data(), hosek.size() * sizeof(float))) return c.fail(RTRT_ERR_IO, "cannot read data/hosek_sky.bin (rtrt_set_data_dir?)");
        if (!c.dHosek) {
            RTRT_CUDA(c, cudaMalloc((void**)&c.dHosek, hosek.size() * sizeof(float)));
            RTRT_CUDA(c, cudaMemcpy(c.dHosek, hosek.data(), hosek.size() * sizeof(float), cudaMemcpyHostToDevice));
        }
        int rc = regenerate_sky(c, hosek);
        if (rc) return rc;
    }
    if (!c.dMaterials || (!c.materialsUser && c.numMaterialsIdx != (int)c.numTris)) {
        int rc = set_materials(c, nullptr, 0, nullptr, 0);
        if (rc) return rc;
    }
    for (int t = 0; t < 2; ++t)
        if (!c.dTexMem[t]) {
            // assets of the reference are missing blobs: constant albedo+AO (mid grey, AO 1) and a flat normal map, roughness 1
            const uint16_t a[4] = {32768, 32768, 32768, 65535}, n[4] = {32768, 32768, 65535, 65535};
            int rc = set_texture(c, t, 1, 1, t == 0 ? a : n);
            if (rc) return rc;
        }
    if (c.width > c.gbufCapW || c.height > c.gbufCapH) {
        // grow: capacity = the largest size seen so far (create the context at maxWidth x maxHeight to avoid this, src/init.cu:55-61)
        RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
        const int capW = c.width > c.gbufCapW ? c.width : c.gbufCapW, capH = c.height > c.gbufCapH ? c.height : c.gbufCapH;
        free_gbuffers(c);
        const size_t px = (size_t)capW * capH;
        const size_t n8 = (size_t)((capW + 7) / 8) * ((capH + 7) / 8), n16 = (size_t)((capW + 15) / 16) * ((capH + 15) / 16);
        struct { void** p; size_t bytes; } plan[] = {
            {&c.dColor, px * 8}, {&c.dNormal, px * 8}, {&c.dAlbedo, px * 8}, {&c.dDepth, px * 2}, {&c.dMotion, px * 4},
            {&c.dAccum, px * 8}, {&c.dHistoryColor, px * 8}, {&c.dHistoryDepth, px * 2}, {&c.dHistoryAlbedo, px * 8},
            {&c.dNoise, n8 * 2}, {&c.dNoise16, n16 * 2}, {&c.dColorTmp, px * 8}, {&c.dGuide, px * 8}};
        for (auto& e : plan) {
            RTRT_CUDA(c, cudaMalloc(e.p, e.bytes));
            RTRT_CUDA(c, cudaMemsetAsync(*e.p, 0, e.bytes, c.stream));
        }
        c.gbufCapW = capW; c.gbufCapH = capH;
    }
    // Dynamic resolution (src/kernel.cu:78-101): within the capacity a new render size only changes the row stride of the buffers
    // written from now on; the history images keep the layout they were written with (accumW x accumH, histW x histH).
    c.gbufW = c.width; c.gbufH = c.height;
    if (c.rowEnd <= c.rowBegin || c.rowEnd > c.height) { c.rowBegin = 0; c.rowEnd = c.height; }
    return RTRT_OK;
}

}  // namespace rtrt
