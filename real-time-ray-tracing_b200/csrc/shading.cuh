// Shading side of the path tracer for sm_100a: texture / environment sampling, BSDFs, light sampling.
// Restates (same formulas, same operation order — radiance has to agree with the oracle to rel-RMSE 1e-3):
//   src/sampler.cuh:293-341,498-584  boundary functors, SampleBicubicSmoothStep(+Lod)
//   src/sky.cuh:33-88                EqualAreaMap / EqualAreaMapCone
//   src/bsdf.cuh                     disk / hemisphere sampling, Lambert, Fresnel, GGX
//   src/light.cuh                    BinarySearch, SampleLight, EnvLight2
//   src/blueNoiseRandGen.h:119-146   Heitz-2019 blue-noise sampler
// Storage differs from the reference on purpose: its cudaArray surfaces (src/init.cu:441-584) become plain
// linear device buffers read through the read-only path (ld.global.nc), which is what B200's L1/L2 serve best.
#pragma once
#include "rtrt_vec.cuh"

namespace rtrt {

constexpr int kSkyW = 512, kSkyH = 256, kSkySize = kSkyW * kSkyH;      // src/kernel.cuh:57-67
constexpr int kSunW = 32, kSunH = 32, kSunSize = kSunW * kSunH;
constexpr int kEnvLightId = 9999, kDefaultLightId = 0x7fffffff;
constexpr int kMaxMips = 11;                                            // src/texture.h:16

enum MatType : uint32_t { LAMBERTIAN_DIFFUSE = 0, PERFECT_REFLECTION = 1, PERFECT_FRESNEL_REFLECTION_REFRACTION = 2,
                          MICROFACET_REFLECTION = 3, EMISSIVE = 4, MAT_SKY = 5 };   // src/kernel.cuh:172-180
struct alignas(16) SurfaceMaterial { Float3 albedo; uint32_t type; Float3 F0; float alpha; };   // src/kernel.cuh:182-196

struct alignas(16) Camera {          // src/kernel.cuh:78-101, also the camera.bin format (src/inputControl.cu:115-150)
    Float3 pos; float pitch; Float3 dir; float focal; Float3 left; float aperture; Float3 up; float yaw;
    Float2 resolution, inversedResolution, fov, tanHalfFov;
    Float3 adjustedLeft; float unused3; Float3 adjustedUp; float unused4; Float3 adjustedFront; float unused5;
    Float3 apertureLeft; float unused6; Float3 apertureUp; float unused7;
};
static_assert(sizeof(Camera) == 176, "Camera layout");

struct HistoryCamera { Mat3 invCamMat; Float3 pos; };                    // src/kernel.cuh:135-155

struct MipTex {                      // src/texture.h:14-25 : 16-bit RGBA mip chain in one linear allocation
    const ushort4* __restrict__ mip[kMaxMips];
    int w[kMaxMips], h[kMaxMips];
    int maxLod;
};

struct EnvMaps {
    const float4* __restrict__ sky; const float* __restrict__ skyCdf;
    const float4* __restrict__ sun; const float* __restrict__ sunCdf;
    Float3 sunDir; float sunAngleCosThetaMax;
};

struct BlueNoiseTables { const uint8_t* __restrict__ sobol; const uint8_t* __restrict__ scramble; const uint8_t* __restrict__ rank; };

// ---- boundary functors src/sampler.cuh:293-341 ----
VD void boundary_repeat(int& x, int& y, int w, int h) {
    if (x >= w) x %= w;
    if (y >= h) y %= h;
    if (x < 0) x = w - (-x) % w;
    if (y < 0) y = h - (-y) % h;
}
VD void boundary_repeatx_clampy(int& x, int& y, int w, int h) {
    if (x >= w) x %= w;
    if (x < 0) x = w - (-x) % w;
    if (y >= h) y = h - 1;
    if (y < 0) y = 0;
}

struct BicubicTaps { int x0, y0; float w[4]; float sum; };
// SampleBicubicSmoothStep src/sampler.cuh:498-541
VD BicubicTaps bicubic_taps(Float2 uv, int w, int h) {
    const Float2 UV = uv * Float2{(float)w, (float)h};
    const float fx0 = floorf(UV.x - 0.5f), fy0 = floorf(UV.y - 0.5f);
    const Float2 f = UV - Float2{fx0 + 0.5f, fy0 + 0.5f};
    const Float2 f2 = f * f, f3_ = f2 * f;
    const Float2 w1{-2.0f * f3_.x + 3.0f * f2.x, -2.0f * f3_.y + 3.0f * f2.y};
    const Float2 w0{1.0f - w1.x, 1.0f - w1.y};
    BicubicTaps t;
    t.x0 = (int)fx0; t.y0 = (int)fy0;
    t.w[0] = w0.x * w0.y; t.w[1] = w1.x * w0.y; t.w[2] = w0.x * w1.y; t.w[3] = w1.x * w1.y;
    t.sum = 0.0f;
#pragma unroll
    for (int i = 0; i < 4; ++i) t.sum += t.w[i];
    return t;
}

// x / 65535.0f for integer x in [0, 65535], correctly rounded: q = x*r, q' = q + r*(x - q*65535) with exact FMAs.
// Verified exhaustively against IEEE division for all 65536 inputs (tools/check_unorm16.c); 3 instructions instead of
// the ~10-instruction division sequence, and the texel loads are the hottest lines of the shading kernels.
VD float unorm16(unsigned short v) {
    const float x = (float)v, r = 1.0f / 65535.0f;
    const float q = __fmul_rn(x, r);
    return __fmaf_rn(__fmaf_rn(-q, 65535.0f, x), r, q);
}
VD Float4 load_ushort4(const ushort4* __restrict__ img, int w, int h, int x, int y) {      // Load2DUshort4 src/sampler.cuh:186-197
    x = clampi(x, 0, w - 1); y = clampi(y, 0, h - 1);                                       // surface clamp mode
    const ushort4 p = __ldg(img + (size_t)y * w + x);
    return Float4{unorm16(p.x), unorm16(p.y), unorm16(p.z), unorm16(p.w)};
}

// BoundaryFuncRepeat (src/sampler.cuh:293-303) followed by the surface clamp, for the tap pair (x0, x0 + 1), with ONE
// integer modulo (a mask when the extent is a power of two) instead of up to four per tap:
//   x in [0,w) -> x ;  x >= w -> x % w ;  x < 0 -> w - (-x) % w  (which can be w: the clamp then gives w - 1)
VD int mod_nonneg(int a, int w) { return ((w & (w - 1)) == 0) ? (a & (w - 1)) : (a % w); }
VD void repeat_pair(int x0, int w, int& a0, int& a1) {
    if (x0 >= 0) {
        const int a = mod_nonneg(x0, w);
        a0 = a; a1 = (a + 1 == w) ? 0 : a + 1;
    } else {
        const int b = mod_nonneg(-x0, w);
        a0 = min(w - b, w - 1);
        a1 = (x0 == -1) ? 0 : min(b == 0 ? 1 : w - b + 1, w - 1);
    }
}

VD Float4 sample_tex_level(const MipTex& m, int lod, Float2 uv) {
    const int w = m.w[lod], h = m.h[lod];
    const BicubicTaps t = bicubic_taps(uv, w, h);
    int xs[2], ys[2];
    repeat_pair(t.x0, w, xs[0], xs[1]);
    repeat_pair(t.y0, h, ys[0], ys[1]);
    Float4 out{0, 0, 0, 0};
#pragma unroll
    for (int i = 0; i < 4; ++i) out = out + load_ushort4(m.mip[lod], w, h, xs[i & 1], ys[i >> 1]) * t.w[i];
    return out / t.sum;
}
// SampleBicubicSmoothStepLod src/sampler.cuh:565-584
VD Float4 sample_tex_lod(const MipTex& m, Float2 uv, float lod) {
    const float low = floorf(lod);
    const float f = lod - low;
    int lo = __float2int_rz(low);                 // cvt.rzi.s32.f32: saturating, NaN -> 0
    int hi = (lo == 0x7fffffff) ? lo : lo + 1;
    lo = clampi(lo, 0, m.maxLod); hi = clampi(hi, 0, m.maxLod);
    const Float4 a = sample_tex_level(m, lo, uv), b = sample_tex_level(m, hi, uv);
    return a * (1.0f - f) + b * f;
}

VD Float3 sample_env(const float4* __restrict__ img, int w, int h, Float2 uv, bool repeatX) {
    const BicubicTaps t = bicubic_taps(uv, w, h);
    Float3 out = f3(0.0f);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int x = t.x0 + (i & 1), y = t.y0 + (i >> 1);
        if (repeatX) boundary_repeatx_clampy(x, y, w, h);
        x = clampi(x, 0, w - 1); y = clampi(y, 0, h - 1);
        const float4 p = __ldg(img + (size_t)y * w + x);
        out = out + Float3{p.x, p.y, p.z} * t.w[i];
    }
    return out / t.sum;
}

// ---- src/sky.cuh:33-88 ----
VD Float3 equal_area_map(float u, float v) {
    const float r = sqrtf(1.0f - v * v);
    const float phi = RTRT_TWO_PI * u;
    return Float3{r * cosf(phi), v, r * sinf(phi)};
}
VD Float2 equal_area_map_dir(Float3 dir) {
    const float u = atan2f(-dir.z, -dir.x) / RTRT_TWO_PI + 0.5f;
    return Float2{u, sel_max(dir.y, 0.05f)};
}
VD Float3 equal_area_map_cone(Float3 sunDir, float u, float v, float cosThetaMax) {
    const float cosTheta = (1.0f - u) + u * cosThetaMax;
    const float sinTheta = sqrtf(1.0f - cosTheta * cosTheta);
    const float phi = v * RTRT_TWO_PI;
    Float3 t, b;
    localize_sample(sunDir, t, b);
    const Mat3 trans = mat3_cols(t, sunDir, b);
    return mul(trans, Float3{cosf(phi) * sinTheta, cosTheta, sinf(phi) * sinTheta});
}
VD bool equal_area_map_cone_inv(Float2& uv, Float3 sunDir, Float3 rayDir, float cosThetaMax) {
    Float3 t, b;
    localize_sample(sunDir, t, b);
    const Mat3 trans = transpose(mat3_cols(t, sunDir, b));
    const Float3 c = mul(trans, rayDir);
    const float cosTheta = c.y;
    if (cosTheta < cosThetaMax) return false;
    const float u = (1.0f - cosTheta) / (1.0f - cosThetaMax);
    const float sinTheta = sqrtf(1.0f - cosTheta * cosTheta);
    if (sinTheta < 1e-5f || (c.x / sinTheta) < -1.0f || (c.x / sinTheta) > 1.0f) return false;
    uv = Float2{u, acosf(c.x / sinTheta) * RTRT_INV_TWO_PI};
    return true;
}

// ---- blue noise src/blueNoiseRandGen.h:119-146 ----
VD float blue_rand(const BlueNoiseTables& bn, int px, int py, int sampleIndex, int dim) {
    px &= 127; py &= 127; sampleIndex &= 255;
    const int tile = (px + py * 128) * 8;
    const int ranked = sampleIndex ^ (int)__ldg(bn.rank + dim + tile);
    int value = (int)__ldg(bn.sobol + dim + ranked * 256);
    value ^= (int)__ldg(bn.scramble + (dim % 8) + tile);
    return (value + 0.5f) / 256.0f;
}

}  // namespace rtrt
