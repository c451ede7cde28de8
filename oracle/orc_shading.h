// ORACLE — TEST INFRASTRUCTURE ONLY (parity pinned against oracle/_ref; see orc_types.h header).
// Shading side of the path tracer: samplers, environment lookups, BSDFs, light sampling, restating
//   src/sampler.cuh:293-341,498-584  boundary functors, SampleBicubicSmoothStep(+Lod)
//   src/sky.cuh:33-88                EqualAreaMap / EqualAreaMapCone
//   src/bsdf.cuh                     disk/hemisphere sampling, Lambert, Fresnel, GGX
//   src/light.cuh                    BinarySearch, SampleLight, EnvLight2, GetLightSource
//   src/blueNoiseRandGen.h:119-146   blue-noise sampler
//   src/kernel.cuh:78-284            Camera / HistoryCamera / SurfaceMaterial / RayState
#pragma once
#include "orc_traverse.h"

namespace orc {

#define ORC_PI_OVER_4 0.7853981633974483096156608458198757210492f
#define ORC_PI_OVER_2 1.5707963267948966192313216916397514420985f
#define ORC_TWO_PI    6.2831853071795864769252867665590057683943f
#define ORC_INV_TWO_PI 0.15915494309f
#define ORC_M_PI 3.1415926535897932384626422832795028841971f   /* src/linearMath.h:14-16: on the reference's platform (MSVC, no
   _USE_MATH_DEFINES) <math.h> does not define M_PI, so the header's own FLOAT constant is the one in effect */
#define SAFE_COSINE_EPSI 1e-5f

constexpr int SKY_WIDTH = 512, SKY_HEIGHT = 256, SKY_SIZE = SKY_WIDTH * SKY_HEIGHT;   // src/kernel.cuh:57-67
constexpr int SUN_WIDTH = 32, SUN_HEIGHT = 32, SUN_SIZE = SUN_WIDTH * SUN_HEIGHT;

struct Camera {            // src/kernel.cuh:78-101 (176 bytes)
    F3 pos; float pitch; F3 dir; float focal; F3 left; float aperture; F3 up; float yaw;
    F2 resolution, inversedResolution, fov, tanHalfFov;
    F3 adjustedLeft; float unused3; F3 adjustedUp; float unused4; F3 adjustedFront; float unused5;
    F3 apertureLeft; float unused6; F3 apertureUp; float unused7;
};
static_assert(sizeof(Camera) == 176, "Camera layout");

// column-major 3x3 (src/linearMath.h:506-539); product uses the compensated inner product
struct Mat3 { float m00, m10, m20, m01, m11, m21, m02, m12, m22; };
static inline Mat3 mat3_cols(F3 a, F3 b, F3 c) { return Mat3{a.x, a.y, a.z, b.x, b.y, b.z, c.x, c.y, c.z}; }
static inline Mat3 transpose(Mat3 m) { Mat3 r = m; r.m01 = m.m10; r.m10 = m.m01; r.m20 = m.m02; r.m02 = m.m20; r.m21 = m.m12; r.m12 = m.m21; return r; }
static inline F3 mul(const Mat3& m, F3 v) {
    return F3{dot(F3{m.m00, m.m01, m.m02}, v), dot(F3{m.m10, m.m11, m.m12}, v), dot(F3{m.m20, m.m21, m.m22}, v)};
}

struct HistoryCamera { Mat3 invCamMat; F3 pos; };        // src/kernel.cuh:135-155
static inline HistoryCamera history_setup(const Camera& c) { return HistoryCamera{transpose(mat3_cols(c.left, c.up, c.dir)), c.pos}; }
static inline F2 WorldToScreenSpace(const HistoryCamera& h, F3 worldPos, F2 tanHalfFov) {
    F3 v = mul(h.invCamMat, worldPos - h.pos);
    F2 sp = F2{v.x, v.y} / v.z;
    F2 ndc = sp / tanHalfFov;
    return F2{0.5f, 0.5f} - ndc * F2{0.5f, 0.5f};
}

enum MatType : u32 { LAMBERTIAN_DIFFUSE = 0, PERFECT_REFLECTION = 1, PERFECT_FRESNEL_REFLECTION_REFRACTION = 2,
                     MICROFACET_REFLECTION = 3, EMISSIVE = 4, MAT_SKY = 5 };     // src/kernel.cuh:172-180
struct SurfaceMaterial { F3 albedo; u32 type; F3 F0; float alpha; };              // src/kernel.cuh:182-196 (32 bytes)

constexpr int ENV_LIGHT_ID = 9999, DEFAULT_LIGHT_ID = 0x7fffffff;                // src/kernel.cuh:37-40 (values only need to differ)

// ---- mip-mapped 16-bit RGBA texture (src/texture.h:14-25) in linear host memory ----
struct Mipmap {
    const uint16_t* mip[11]; int w[11], h[11]; int maxLod;
};
static inline F4 f4(float a, float b, float c, float d) { return F4{a, b, c, d}; }
static inline F4 operator*(F4 a, float s) { return F4{a.x * s, a.y * s, a.z * s, a.w * s}; }
static inline F4 operator+(F4 a, F4 b) { return F4{a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
static inline F4 operator/(F4 a, float s) { return F4{a.x / s, a.y / s, a.z / s, a.w / s}; }

// boundary functors src/sampler.cuh:293-341
static inline void boundary_repeat(int& x, int& y, int w, int h) {
    if (x >= w) x %= w;
    if (y >= h) y %= h;
    if (x < 0) x = w - (-x) % w;
    if (y < 0) y = h - (-y) % h;
}
static inline void boundary_repeatx_clampy(int& x, int& y, int w, int h) {
    if (x >= w) x %= w;
    if (x < 0) x = w - (-x) % w;
    if (y >= h) y = h - 1;
    if (y < 0) y = 0;
}
// surf2Dread with cudaBoundaryModeClamp clamps out-of-range coordinates (BoundaryFuncRepeat can return x == w)
static inline int clampi(int a, int lo, int hi) { return a < lo ? lo : a > hi ? hi : a; }

// Load2DUshort4 src/sampler.cuh:186-197
static inline F4 load_ushort4(const uint16_t* img, int w, int h, int x, int y) {
    x = clampi(x, 0, w - 1); y = clampi(y, 0, h - 1);
    const uint16_t* p = img + 4 * ((size_t)y * w + x);
    return F4{p[0] / 65535.0f, p[1] / 65535.0f, p[2] / 65535.0f, p[3] / 65535.0f};
}
static inline F3 load_float4_xyz(const float* img, int w, int h, int x, int y) {
    x = clampi(x, 0, w - 1); y = clampi(y, 0, h - 1);
    const float* p = img + 4 * ((size_t)y * w + x);
    return F3{p[0], p[1], p[2]};
}

// SampleBicubicSmoothStep src/sampler.cuh:498-541 — 4 taps, smoothstep weights, normalised
struct BicubicTaps { int x0, y0, x1, y1; float w[4]; float sum; };
static inline BicubicTaps bicubic_taps(F2 uv, int w, int h) {
    F2 UV = uv * F2{(float)w, (float)h};
    F2 tc = F2{floorf(UV.x - 0.5f) + 0.5f, floorf(UV.y - 0.5f) + 0.5f};
    F2 f = UV - tc;
    F2 f2 = f * f, f3 = f2 * f;
    F2 w1 = F2{-2.0f * f3.x + 3.0f * f2.x, -2.0f * f3.y + 3.0f * f2.y};
    F2 w0 = F2{1.0f - w1.x, 1.0f - w1.y};
    BicubicTaps t;
    t.x0 = (int)floorf(UV.x - 0.5f); t.y0 = (int)floorf(UV.y - 0.5f); t.x1 = t.x0 + 1; t.y1 = t.y0 + 1;
    t.w[0] = w0.x * w0.y; t.w[1] = w1.x * w0.y; t.w[2] = w0.x * w1.y; t.w[3] = w1.x * w1.y;
    t.sum = 0; for (int i = 0; i < 4; ++i) t.sum += t.w[i];
    return t;
}
static inline F4 sample_tex_level(const Mipmap& m, int lod, F2 uv) {              // BoundaryFuncRepeat + Load2DFuncUshort4<Float4>
    int w = m.w[lod], h = m.h[lod];
    BicubicTaps t = bicubic_taps(uv, w, h);
    int xs[4] = {t.x0, t.x1, t.x0, t.x1}, ys[4] = {t.y0, t.y0, t.y1, t.y1};
    F4 out{0, 0, 0, 0};
    for (int i = 0; i < 4; ++i) {
        int x = xs[i], y = ys[i];
        boundary_repeat(x, y, w, h);
        out = out + load_ushort4(m.mip[lod], w, h, x, y) * t.w[i];
    }
    return out / t.sum;
}
// SampleBicubicSmoothStepLod src/sampler.cuh:565-584
static inline F4 sample_tex_lod(const Mipmap& m, F2 uv, float lod) {
    float low = floorf(lod);
    float f = lod - low;
    int lo = (int)low, hi = lo + 1;     // (int) of -inf/NaN: CUDA cvt.rzi.s32.f32 saturates / gives 0; clamp below covers it
    if (!(low == low)) lo = 0, hi = 1;
    else if (low <= -2147483648.0f) lo = INT32_MIN, hi = INT32_MIN + 1;
    else if (low >= 2147483647.0f) lo = INT32_MAX, hi = INT32_MAX;
    lo = clampi(lo, 0, m.maxLod); hi = clampi(hi, 0, m.maxLod);
    F4 a = sample_tex_level(m, lo, uv), b = sample_tex_level(m, hi, uv);
    return a * (1.0f - f) + b * f;
}
// float4 environment maps (sky: RepeatX/ClampY; sun: default boundary + surface clamp)
static inline F3 sample_env(const float* img, int w, int h, F2 uv, bool repeatX) {
    BicubicTaps t = bicubic_taps(uv, w, h);
    int xs[4] = {t.x0, t.x1, t.x0, t.x1}, ys[4] = {t.y0, t.y0, t.y1, t.y1};
    F3 out = f3(0.0f);
    for (int i = 0; i < 4; ++i) {
        int x = xs[i], y = ys[i];
        if (repeatX) boundary_repeatx_clampy(x, y, w, h);
        out = out + load_float4_xyz(img, w, h, x, y) * t.w[i];
    }
    return out / t.sum;
}

// ---- src/linearMath.h:725-738 LocalizeSample ----
static inline void LocalizeSample(F3 n, F3& u, F3& v) {
    F3 w = f3(1, 0, 0);
    if (fabsf(n.x) > 0.707f) w = f3(0, 1, 0);
    u = cross(n, w);
    v = cross(n, u);
}

// ---- src/sky.cuh:33-88 ----
static inline F3 EqualAreaMap(float u, float v) {
    float z = v;
    float r = sqrtf(1.0f - v * v);
    float phi = ORC_TWO_PI * u;
    return F3{r * cosf(phi), z, r * sinf(phi)};
}
static inline F2 EqualAreaMapDir(F3 dir) {
    float u = atan2f(-dir.z, -dir.x) / ORC_TWO_PI + 0.5f;
    float v = rmax(dir.y, 0.05f);
    return F2{u, v};
}
static inline F3 EqualAreaMapCone(F3 sunDir, float u, float v, float cosThetaMax) {
    float cosTheta = (1.0f - u) + u * cosThetaMax;
    float sinTheta = sqrtf(1.0f - cosTheta * cosTheta);
    float phi = v * ORC_TWO_PI;
    F3 t, b;
    LocalizeSample(sunDir, t, b);
    Mat3 trans = mat3_cols(t, sunDir, b);
    F3 coords = F3{cosf(phi) * sinTheta, cosTheta, sinf(phi) * sinTheta};
    return mul(trans, coords);
}
static inline bool EqualAreaMapConeInv(F2& uv, F3 sunDir, F3 rayDir, float cosThetaMax) {
    F3 t, b;
    LocalizeSample(sunDir, t, b);
    Mat3 trans = transpose(mat3_cols(t, sunDir, b));
    F3 coords = mul(trans, rayDir);
    float cosTheta = coords.y;
    if (cosTheta < cosThetaMax) return false;
    float u = (1.0f - cosTheta) / (1.0f - cosThetaMax);
    float sinTheta = sqrtf(1.0f - cosTheta * cosTheta);
    if (sinTheta < 1e-5f || (coords.x / sinTheta) < -1.0f || (coords.x / sinTheta) > 1.0f) return false;
    float v = acosf(coords.x / sinTheta) * ORC_INV_TWO_PI;
    uv = F2{u, v};
    return true;
}

// (continued in orc_shading_bsdf.h)
}  // namespace orc
