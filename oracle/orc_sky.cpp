// ORACLE — TEST INFRASTRUCTURE ONLY (parity pinned against oracle/_ref; see orc_types.h header).
// Environment inputs of the path tracer (regenerated only when the sky parameters change), restating
//   src/kernel.cu:121-124,289-308   sun direction, regeneration sequence, sunAngleCosThetaMax
//   src/sky.cuh:91-319              Hosek-Wilkie sky radiance (10 bands), solar disc radiance, Sky / SkySun kernels
//   src/color.h:19-30               XyzToRgbSrgb
//   src/scan.cuh:235-251            CpuScan (the reference's own CPU scan; postfix = inclusive)
#include "orc_shading.h"
#include <vector>

namespace orc {

struct SkyParams { float timeOfDay, sunAxisAngle, skyScalar, sunScalar, sunAngle; };   // src/settingParams.h:26-46
struct SkyData { const float* skyDataSets; const float* skyDataSetsRad; const float* solarDatasets; const float* limbDarkening; };

static const float spectrumCieX[10] = {2.372527e-02f, 1.955480e+00f, 1.074553e+01f, 5.056697e+00f, 4.698190e+00f, 2.391135e+01f, 3.798705e+01f, 1.929414e+01f, 2.970610e+00f, 2.092986e-01f};
static const float spectrumCieY[10] = {6.813859e-04f, 6.771017e-02f, 1.171193e+00f, 6.997765e+00f, 2.666710e+01f, 3.758372e+01f, 2.503930e+01f, 8.150395e+00f, 1.098635e+00f, 7.563256e-02f};
static const float spectrumCieZ[10] = {1.119121e-01f, 9.441195e+00f, 5.597921e+01f, 3.589996e+01f, 5.070894e+00f, 3.523189e-01f, 3.422707e-02f, 2.539118e-03f, 7.836666e-06f, 0.000000e+00f};

static F3 XyzToRgbSrgb(F3 c) {      // src/color.h:19-30 (row-major constructor arguments, compensated products)
    return F3{dot(F3{3.2404542f, -1.5371385f, -0.4985314f}, c), dot(F3{-0.9692660f, 1.8760108f, 0.0415560f}, c),
              dot(F3{0.0556434f, -0.2040259f, 1.0572252f}, c)};
}

// src/sky.cuh:91-112
static float GetFittingData(const float* m, float e, int i) {
    return (powf(1.0f - e, 5.0f) * m[i] + 5.0f * powf(1.0f - e, 4.0f) * e * m[i + 9] + 10.0f * powf(1.0f - e, 3.0f) * powf(e, 2.0f) * m[i + 18] +
            10.0f * powf(1.0f - e, 2.0f) * powf(e, 3.0f) * m[i + 27] + 5.0f * (1.0f - e) * powf(e, 4.0f) * m[i + 36] + powf(e, 5.0f) * m[i + 45]);
}
static float GetFittingData2(const float* m, float e) {
    return (powf(1.0f - e, 5.0f) * m[0] + 5.0f * powf(1.0f - e, 4.0f) * e * m[1] + 10.0f * powf(1.0f - e, 3.0f) * powf(e, 2.0f) * m[2] +
            10.0f * powf(1.0f - e, 2.0f) * powf(e, 3.0f) * m[3] + 5.0f * (1.0f - e) * powf(e, 4.0f) * m[4] + powf(e, 5.0f) * m[5]);
}
// UpdateSkyState src/sky.cuh:126-148 (host code in the reference as well)
static void UpdateSkyState(const SkyData& d, F3 sunDir, float* cfg /*90*/, float* rad /*10*/) {
    float elevation = acosf(sunDir.y);
    float se = powf(elevation / (ORC_M_PI / 2.0f), (1.0f / 3.0f));
    for (int ch = 0; ch < 10; ++ch) {
        for (int i = 0; i < 9; ++i) cfg[ch * 9 + i] = GetFittingData(d.skyDataSets + ch * 54, se, i);
        rad[ch] = GetFittingData2(d.skyDataSetsRad + ch * 6, se);
    }
}

// src/sky.cuh:166-197
static F3 GetSkyRadiance(F3 raydir, F3 sunDir, const float* cfg, const float* rad) {
    float theta = acosf(raydir.y);
    float gamma = acosf(clampf(dot(raydir, sunDir), -1, 1));
    F3 xyz = f3(0.0f);
    for (int ch = 0; ch < 10; ++ch) {
        const float* c = cfg + ch * 9;
        const float expM = expf(c[4] * gamma);
        const float rayM = cosf(gamma) * cosf(gamma);
        const float mieM = (1.0f + cosf(gamma) * cosf(gamma)) / powf((1.0f + c[8] * c[8] - 2.0f * c[8] * cosf(gamma)), 1.5f);
        const float zenith = sqrtf(cosf(theta));
        float ri = (1.0f + c[0] * expf(c[1] / (cosf(theta) + 0.01f))) * (c[2] + c[3] * expM + c[5] * rayM + c[6] * mieM + c[7] * zenith);
        float radiance = ri * rad[ch];
        xyz = xyz + radiance * F3{spectrumCieX[ch], spectrumCieY[ch], spectrumCieZ[ch]};
    }
    return XyzToRgbSrgb(xyz);
}

// src/sky.cuh:199-277
static F3 GetSunRadiance(F3 raydir, F3 sunDir, const SkyParams& p, const SkyData& d) {
    float gamma = acosf(clampf(dot(raydir, sunDir), -1, 1));
    float elevation = (ORC_M_PI / 2.0f) - acosf(sunDir.y);
    const float solarRadius = p.sunAngle * ORC_M_PI / 180.0f / 2.0f;
    const float scale = 1.0f / ((p.sunAngle / 0.51f) * (p.sunAngle / 0.51f));
    F3 xyz = f3(0.0f);
    float srs = sinf(solarRadius);
    float ar2 = 1.0f / (srs * srs);
    float sg = sinf(gamma);
    float sc2 = 1.0f - ar2 * sg * sg;
    if (sc2 < 0.0f) sc2 = 0.0f;
    float sampleCosine = sqrtf(sc2);
    if (sampleCosine == 0.0f) return f3(0.0f);
    for (int ch = 0; ch < 10; ++ch) {
        const int pieces = 45, order = 4;
        int pos = (int)(powf((float)(2.0 * elevation / ORC_M_PI), (float)(1.0 / 3.0)) * pieces);
        if (pos > 44) pos = 44;
        const float break_x = (float)(powf(((float)pos / (float)pieces), 3.0f) * ((double)ORC_M_PI * 0.5));
        const float* coefs = d.solarDatasets + ch * 180 + (order * (pos + 1) - 1);
        float res = 0.0f;
        const float x = elevation - break_x;
        float x_exp = 1.0f;
        for (int i = 0; i < order; ++i) { res += x_exp * *coefs--; x_exp *= x; }
        float direct = res;
        const float* ld = d.limbDarkening + ch * 6;
        float dark = ld[0] + ld[1] * sampleCosine + ld[2] * powf(sampleCosine, 2.0f) + ld[3] * powf(sampleCosine, 3.0f) +
                     ld[4] * powf(sampleCosine, 4.0f) + ld[5] * powf(sampleCosine, 5.0f);
        direct *= dark * scale;
        xyz = xyz + direct * F3{spectrumCieX[ch], spectrumCieY[ch], spectrumCieZ[ch]};
    }
    return XyzToRgbSrgb(xyz);
}

static inline F3 EqualAreaMapS(float u, float v) {
    float r = sqrtf(1.0f - v * v), phi = ORC_TWO_PI * u;
    return F3{r * cosf(phi), v, r * sinf(phi)};
}

}  // namespace orc

using namespace orc;

extern "C" {

// src/kernel.cu:121-124 : axis, angle, rotate3f(axis, angle, cross(up, axis)).normalized()
void orc_sun_dir(float timeOfDay, float sunAxisAngleDeg, float* out3) {
    const float d2r = 0.01745329251f;                                         // Pi_over_180
    F3 axis = normalize(F3{0.0f, cosf(sunAxisAngleDeg * d2r), sinf(sunAxisAngleDeg * d2r)});
    float angle = fmodf(timeOfDay * ORC_M_PI, ORC_TWO_PI);
    F3 v = cross(f3(0, 1, 0), axis);
    // q * v * q^-1 with q = (axis.normalized()*sin(a/2), cos(a/2))   (src/linearMath.h:696-715)
    F3 qv = normalize(axis) * sinf(angle / 2); float qw = cosf(angle / 2);
    // p = q * (v,0)
    F3 pv = qw * v + 0.0f * qv + cross(qv, v); float pw = qw * 0.0f - dot(qv, v);
    // r = p * conj(q)
    F3 cv = -qv; float cw = qw;
    F3 rv = pw * cv + cw * pv + cross(pv, cv);
    F3 s = normalize(rv);
    out3[0] = s.x; out3[1] = s.y; out3[2] = s.z;
}

// data: skyDataSets[540] | skyDataSetsRad[60] | solarDatasets[1800] | limbDarkening[60]
// outputs: sky float4[512*256], skyCdf[131072], sun float4[32*32], sunCdf[1024], cosThetaMax
int orc_sky_generate(const float* params5, const float* sunDir3, const float* data,
                     float* sky, float* skyCdf, float* sun, float* sunCdf, float* sunAngleCosThetaMax) {
    SkyParams p{params5[0], params5[1], params5[2], params5[3], params5[4]};
    p.sunScalar = rmax(p.sunScalar, 0.00001f); p.skyScalar = rmax(p.skyScalar, 0.00001f); p.sunAngle = rmax(p.sunAngle, 0.51f);   // src/kernel.cu:291-293
    SkyData d{data, data + 540, data + 600, data + 2400};
    F3 sunDir{sunDir3[0], sunDir3[1], sunDir3[2]};
    float cfg[90], rad[10];
    UpdateSkyState(d, sunDir, cfg, rad);
    std::vector<float> pdf(SKY_SIZE);
    for (int y = 0; y < SKY_HEIGHT; ++y)
        for (int x = 0; x < SKY_WIDTH; ++x) {                                   // Sky kernel src/sky.cuh:280-299
            float u = ((float)x + 0.5f) / SKY_WIDTH, v = ((float)y + 0.5f) / SKY_HEIGHT;
            F3 c = GetSkyRadiance(EqualAreaMapS(u, v), sunDir, cfg, rad) * p.skyScalar;
            c = max3f(c, f3(0.0f));
            int i = SKY_WIDTH * y + x;
            sky[4 * i] = c.x; sky[4 * i + 1] = c.y; sky[4 * i + 2] = c.z; sky[4 * i + 3] = 0;
            pdf[i] = dot(c, F3{0.3f, 0.6f, 0.1f});
        }
    float acc = 0; for (int i = 0; i < SKY_SIZE; ++i) { acc += pdf[i]; skyCdf[i] = acc; }   // CpuScan(postfix=1)
    float cosMax = cosf(p.sunAngle * ORC_M_PI / 180.0f / 2.0f);
    std::vector<float> spdf(SUN_SIZE);
    for (int y = 0; y < SUN_HEIGHT; ++y)
        for (int x = 0; x < SUN_WIDTH; ++x) {                                   // SkySun kernel src/sky.cuh:301-319
            float u = ((float)x + 0.5f) / SUN_WIDTH, v = ((float)y + 0.5f) / SUN_HEIGHT;
            F3 dir = EqualAreaMapCone(sunDir, u, v, cosMax);
            F3 c = GetSunRadiance(dir, sunDir, p, d) * p.sunScalar;
            c = max3f(c, f3(0.0f));
            int i = SUN_WIDTH * y + x;
            sun[4 * i] = c.x; sun[4 * i + 1] = c.y; sun[4 * i + 2] = c.z; sun[4 * i + 3] = 0;
            spdf[i] = dot(c, F3{0.3f, 0.6f, 0.1f});
        }
    acc = 0; for (int i = 0; i < SUN_SIZE; ++i) { acc += spdf[i]; sunCdf[i] = acc; }
    *sunAngleCosThetaMax = cosMax;                                              // src/kernel.cu:303-305
    return 0;
}

}  // extern "C"
