// ORACLE — TEST INFRASTRUCTURE ONLY (parity pinned against oracle/_ref; see orc_types.h header).
// Stage 3: the reference's temporal + spatial denoiser with PING-PONG semantics (the reference filters in place,
// so its own output is race dependent — src/temporalDenoising.cuh:556,607), restating
//   src/denoising.cu:48-186           pass sequence and toggles
//   src/temporalDenoising.cuh:10-21   RgbToYcocg / YcocgToRgb
//   :33-102   CalculateTileNoiseLevel, TileNoiseLevel8x8to16x16     :317-492  SpatialFilter7x7
//   :495-608  SpatialFilterGlobal5x5<3|6|12>                        :610-893  TemporalFilter
//   :896-1111 TemporalFilter2     :1127-1139 ApplyAlbedo            :142-170  history copies
//   src/gaussian.cuh:12-43            Gaussian tables
// Every pass reads half-precision buffers and stores half-precision results exactly where the reference does.
#include "orc_shading.h"
#include <vector>
#include <thread>
#include <functional>

namespace orc {

typedef uint16_t h16;
static inline float h2f(h16 b) { _Float16 h; std::memcpy(&h, &b, 2); return (float)h; }
static inline h16 f2h(float f) { _Float16 h = (_Float16)f; h16 b; std::memcpy(&b, &h, 2); return b; }   // round-to-nearest-even

struct Px4 { h16 x, y, z, w; };     // half3 + (u16 mask | half)

struct DenoiseParams {
    float local_n, local_d, local_m, large_n, large_d, large_m, temp_n, temp_d, temp_m, thr_local, thr_large;
};

static const float G3[9] = {0.0578968, 0.0921378, 0.0584323, 0.0921378, 0.146629, 0.09299, 0.0584322, 0.0929898, 0.0589727};
static const float G5[25] = {0.00360466, 0.0144464, 0.0229902, 0.01458, 0.0036719, 0.0144464, 0.0578968, 0.0921378, 0.0584323, 0.0147159,
                             0.0229902, 0.0921378, 0.146629, 0.09299, 0.023419, 0.01458, 0.0584322, 0.0929898, 0.0589727, 0.014852,
                             0.00367191, 0.0147158, 0.0234191, 0.0148519, 0.0037404};
static const float G7[49] = {3.47404e-05, 0.000353875, 0.00141822, 0.00225698, 0.00143134, 0.000360475, 3.57221e-05,
                             0.000353875, 0.00360466, 0.0144464, 0.0229902, 0.01458, 0.0036719, 0.000363875,
                             0.00141822, 0.0144464, 0.0578968, 0.0921378, 0.0584323, 0.0147159, 0.0014583,
                             0.00225698, 0.0229902, 0.0921378, 0.146629, 0.09299, 0.023419, 0.00232076,
                             0.00143134, 0.01458, 0.0584322, 0.0929898, 0.0589727, 0.014852, 0.00147179,
                             0.000360475, 0.00367191, 0.0147158, 0.0234191, 0.0148519, 0.0037404, 0.000370662,
                             3.57221e-05, 0.000363875, 0.0014583, 0.00232075, 0.00147179, 0.000370662, 3.67315e-05};

static inline F3 RgbToYcocg(F3 rgb) { float t1 = rgb.x + rgb.z, t2 = rgb.y * 2.0f; return F3{t1 + t2, (rgb.x - rgb.z) * 2.0f, t2 - t1}; }
static inline F3 YcocgToRgb(F3 y) { float t = y.x - y.z; return F3{t + y.y, y.x + y.z, t - y.y} * 0.25f; }
static inline bool nan3(F3 v) { return v.x != v.x || v.y != v.y || v.z != v.z; }

struct Img {                      // clamp-addressed views (surface reads use cudaBoundaryModeClamp)
    int W, H;
    const Px4* c;
    Px4 at(int x, int y) const { x = clampi(x, 0, W - 1); y = clampi(y, 0, H - 1); return c[(size_t)y * W + x]; }
};
static inline F3 rgb(Px4 p) { return F3{h2f(p.x), h2f(p.y), h2f(p.z)}; }
static inline Px4 pack(F3 c, h16 w) { return Px4{f2h(c.x), f2h(c.y), f2h(c.z), w}; }

struct Frame {
    int W, H;
    const Px4* normal; const h16* depth; const h16* motion; const Px4* albedo;
    F3 nrm(int x, int y) const { x = clampi(x, 0, W - 1); y = clampi(y, 0, H - 1); return rgb(normal[(size_t)y * W + x]); }
    float dep(int x, int y) const { x = clampi(x, 0, W - 1); y = clampi(y, 0, H - 1); return h2f(depth[(size_t)y * W + x]); }
};

static void parallel_rows(int H, int threads, const std::function<void(int)>& fn) {
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t) pool.emplace_back([&, t]() { for (int y = t; y < H; y += threads) fn(y); });
    for (auto& th : pool) th.join();
}

// SampleBicubicSmoothStep<Load2DFuncHalf3Ushort1> with the default boundary functor (src/sampler.cuh:498-541)
static F3 sample_history(const Img& img, F2 uv) {
    BicubicTaps t = bicubic_taps(uv, img.W, img.H);
    int xs[4] = {t.x0, t.x1, t.x0, t.x1}, ys[4] = {t.y0, t.y0, t.y1, t.y1};
    F3 out = f3(0.0f);
    for (int i = 0; i < 4; ++i) out = out + rgb(img.at(xs[i], ys[i])) * t.w[i];
    return out / t.sum;
}

// shared tail of TemporalFilter / TemporalFilter2 (:861-889, :1079-1107)
static F3 temporal_blend(F3 colorValue, F3 colorHistory, float lumaHistory, float lumaMin, float lumaMax, float lumaCurrent, float base) {
    float blend = base;
    blend *= 0.2f + 0.8f * clampf(0.5f * rmin(fabsf(lumaHistory - lumaMin), fabsf(lumaHistory - lumaMax)) / rmax(rmax(lumaHistory, lumaCurrent), 1e-4f));
    float wA = blend * rmax(0.0001f, 1.0f / (lumaCurrent + 4.0f));
    float wB = (1.0f - blend) * rmax(0.0001f, 1.0f / (lumaHistory + 4.0f));
    float ws = SafeDivide(1.0f, wA + wB);
    wA *= ws; wB *= ws;
    return colorValue * wA + colorHistory * wB;
}

// TemporalFilter src/temporalDenoising.cuh:610-893
static void temporal_filter(const Frame& f, const Img& in, const Img& accum, Px4* out, const DenoiseParams& p, int threads) {
    parallel_rows(f.H, threads, [&](int y) {
        for (int x = 0; x < f.W; ++x) {
            Px4 cpx = in.at(x, y);
            out[(size_t)y * f.W + x] = cpx;
            F3 colorValue = rgb(cpx); float depthValue = f.dep(x, y); F3 normalValue = f.nrm(x, y); h16 maskValue = cpx.w;
            if (depthValue >= RayMax) continue;
            F3 nMax = RgbToYcocg(colorValue), nMin = nMax;
            F3 filtered = f3(0.0f); float wsum = 0;
            for (int j = 0; j < 9; ++j) {
                int xo = j % 3, yo = j / 3;
                Px4 q = in.at(x + xo - 1, y + yo - 1);
                F3 color = rgb(q); float depth = f.dep(x + xo - 1, y + yo - 1); F3 normal = f.nrm(x + xo - 1, y + yo - 1);
                float w = 1.0f;
                w *= powf(rmax(dot(normalValue, normal), 0.0f), p.temp_n);
                float dd = (depthValue - depth) / p.temp_d;
                w *= expf(-0.5f * dd * dd);
                w *= (maskValue != q.w) ? 1.0f / p.temp_m : 1.0f;
                w *= G3[xo + yo * 3];
                filtered = filtered + color * w; wsum += w;
                F3 nc = RgbToYcocg(color);
                nMax = max3f(nMax, nc); nMin = min3f(nMin, nc);
            }
            if (wsum > 0) filtered = filtered / wsum; else filtered = f3(0.0f);
            if (nan3(filtered)) filtered = f3(0.0f);
            const h16* mv = f.motion + 2 * ((size_t)y * f.W + x);
            F2 motionVec = F2{h2f(mv[0]), h2f(mv[1])} - 0.5f;
            F2 uv = (F2{(float)x, (float)y} + 0.5f) * F2{1.0f / (float)f.W, 1.0f / (float)f.H};
            F2 huv = uv + motionVec;
            if (huv.x < 0 || huv.y < 0 || huv.x > 1.0 || huv.y > 1.0) { out[(size_t)y * f.W + x] = pack(filtered, maskValue); continue; }
            F3 hist = sample_history(accum, huv);
            F3 hy = RgbToYcocg(hist);
            hy = F3{clampf(hy.x, nMin.x, nMax.x), clampf(hy.y, nMin.y, nMax.y), clampf(hy.z, nMin.z, nMax.z)};
            hist = YcocgToRgb(hy);
            float lumaMin = nMin.x, lumaMax = nMax.x, lumaCurrent = RgbToYcocg(colorValue).x;
            float discard = 0;
            int hx = (int)floorf(huv.x * accum.W), hyy = (int)floorf(huv.y * accum.H);
            for (int i = 0; i < 4; ++i) discard += (maskValue != accum.at(hx + i % 2, hyy + i / 2).w);
            discard /= 4.0f;
            hist = hist * (1.0f - discard) + filtered * discard;
            float lumaHistory = RgbToYcocg(hist).x;
            if (nan3(hist)) hist = f3(0.0f);
            F3 o = temporal_blend(colorValue, hist, lumaHistory, lumaMin, lumaMax, lumaCurrent, 1.0f / 8.0f);
            if (nan3(o)) o = f3(0.0f);
            out[(size_t)y * f.W + x] = pack(o, maskValue);
        }
    });
}

// shfl_down tree of WarpReduceSum (:23-31): lane 0 ends with this association order
static float tree32(float* a) { for (int o = 16; o > 0; o >>= 1) for (int i = 0; i < o; ++i) a[i] += a[i + o]; return a[0]; }

// CalculateTileNoiseLevel (:33-91) + TileNoiseLevel8x8to16x16 (:93-102)
static void noise_levels(const Frame& f, const Img& in, h16* n8, h16* n16) {
    const int gw = (f.W + 7) / 8, gh = (f.H + 7) / 8, g16w = (f.W + 15) / 16, g16h = (f.H + 15) / 16;
    for (int by = 0; by < gh; ++by)
        for (int bx = 0; bx < gw; ++bx) {
            float l1[32], l12[32], l2[32], l22[32], b1[32], b2[32];
            for (int ty = 0; ty < 4; ++ty)
                for (int tx = 0; tx < 8; ++tx) {
                    int lane = tx + 8 * ty, x = tx + bx * 8, y = (ty + by * 4) * 2;
                    F3 c1 = rgb(in.at(x, y)), c2 = rgb(in.at(x, y + 1));
                    b1[lane] = f.dep(x, y) >= RayMax ? 1.0f : 0.0f; b2[lane] = f.dep(x, y + 1) >= RayMax ? 1.0f : 0.0f;
                    l1[lane] = getmax(c1); l12[lane] = l1[lane] * l1[lane];
                    l2[lane] = getmax(c2); l22[lane] = l2[lane] * l2[lane];
                }
            float bg = tree32(b1) + tree32(b2);       // integer sums in the reference (uint): exact either way
            float s1 = tree32(l1), s12 = tree32(l12), s2 = tree32(l2), s22 = tree32(l22);
            float notSky = 1.0f - bg / 64.0f;
            float ave = (s1 + s2) / 64.0f, aveSq = ave * ave, sqAve = (s12 + s22) / 64.0f;
            float var = rmax(1e-20f, sqAve - aveSq);
            float noise = var / rmax(aveSq, 1e-20f);
            noise *= notSky;
            n8[(size_t)by * gw + bx] = f2h(noise);
        }
    auto at8 = [&](int x, int y) { x = clampi(x, 0, gw - 1); y = clampi(y, 0, gh - 1); return h2f(n8[(size_t)y * gw + x]); };
    for (int y = 0; y < g16h; ++y)
        for (int x = 0; x < g16w; ++x)
            n16[(size_t)y * g16w + x] = f2h((at8(2 * x, 2 * y) + at8(2 * x + 1, 2 * y) + at8(2 * x, 2 * y + 1) + at8(2 * x + 1, 2 * y + 1)) / 4);
}

// SpatialFilter7x7 (:317-492): 24 of the 49 taps, start tap = frameNum % 2
static void spatial7x7(const Frame& f, const Img& in, Px4* out, const h16* n16, int frameNum, const DenoiseParams& p, int threads) {
    const int g16w = (f.W + 15) / 16;
    parallel_rows(f.H, threads, [&](int y) {
        for (int x = 0; x < f.W; ++x) {
            Px4 cpx = in.at(x, y);
            out[(size_t)y * f.W + x] = cpx;
            if (h2f(n16[(size_t)(y / 16) * g16w + x / 16]) < p.thr_local) continue;
            F3 normalValue = f.nrm(x, y); float depthValue = f.dep(x, y); h16 maskValue = cpx.w;
            if (depthValue >= RayMax) continue;
            F3 sumC = f3(0.0f); float sumW = 0;
            int j = frameNum % 2;
            for (int i = 0; i < 24; ++i) {
                int xo = j % 7, yo = j / 7; j += 2;
                int sx = x + xo - 3, sy = y + yo - 3;
                Px4 q = in.at(sx, sy);
                F3 color = rgb(q); float depth = f.dep(sx, sy); F3 normal = f.nrm(sx, sy);
                float w = 1.0f;
                w *= powf(rmax(dot(normalValue, normal), 0.0001f), p.local_n);
                float dd = (depthValue - depth) / p.local_d;
                w *= expf(-0.5f * dd * dd);
                w *= (maskValue != q.w) ? 1.0f / p.local_m : 1.0f;
                w *= G7[xo + yo * 7];
                sumC = sumC + color * w; sumW += w;
            }
            if (nan3(sumC)) sumC = f3(0.0f);
            if (sumW != sumW) sumW = 0;
            F3 fin = (sumW == 0) ? f3(0.0f) : sumC / sumW;
            if (nan3(fin)) fin = f3(0.0f);
            out[(size_t)y * f.W + x] = pack(fin, maskValue);
        }
    });
}

// SpatialFilterGlobal5x5<S> (:495-608)
static void spatial5x5(const Frame& f, const Img& in, Px4* out, const h16* n16, int S, const DenoiseParams& p, int threads) {
    const int g16w = (f.W + 15) / 16;
    parallel_rows(f.H, threads, [&](int y) {
        for (int x = 0; x < f.W; ++x) {
            Px4 cpx = in.at(x, y);
            out[(size_t)y * f.W + x] = cpx;
            if (h2f(n16[(size_t)(y / 16) * g16w + x / 16]) < p.thr_large) continue;
            F3 normalValue = f.nrm(x, y); float depthValue = f.dep(x, y); h16 maskValue = cpx.w;
            if (depthValue >= 10e9f) continue;
            F3 sumC = f3(0.0f); float sumW = 0;
            for (int k = 0; k < 25; ++k) {
                int i = k % 5, j = k / 5;
                int sx = x + (i - 2) * S, sy = y + (j - 2) * S;
                Px4 q = in.at(sx, sy);
                F3 color = rgb(q); float depth = f.dep(sx, sy); F3 normal = f.nrm(sx, sy);
                float w = 1.0f;
                w *= powf(rmax(dot(normalValue, normal), 0.0f), p.large_n);
                float dd = (depthValue - depth) / p.large_d;
                w *= expf(-0.5f * dd * dd);
                w *= (maskValue != q.w) ? 1.0f / p.large_m : 1.0f;
                w *= G5[i + j * 5];
                if (nan3(color)) { color = f3(0.0f); w = 0; }
                sumC = sumC + color * w; sumW += w;
            }
            if (nan3(sumC)) sumC = f3(0.0f);
            if (sumW != sumW) sumW = 0;
            F3 fin = (sumW == 0) ? f3(0.0f) : sumC / sumW;
            if (nan3(fin)) fin = f3(0.0f);
            out[(size_t)y * f.W + x] = pack(fin, maskValue);
        }
    });
}

// ApplyAlbedo (:1127-1139): colour *= albedo, the mask channel becomes half(1.0)
static void apply_albedo(const Frame& f, const Img& in, Px4* out, int threads) {
    parallel_rows(f.H, threads, [&](int y) {
        for (int x = 0; x < f.W; ++x) {
            F3 c = rgb(in.at(x, y)) * rgb(f.albedo[(size_t)y * f.W + x]);
            out[(size_t)y * f.W + x] = Px4{f2h(c.x), f2h(c.y), f2h(c.z), f2h(1.0f)};
        }
    });
}

// TemporalFilter2 (:896-1111)
static void temporal_filter2(const Frame& f, const Img& in, const Img& hist, Px4* out, int threads) {
    parallel_rows(f.H, threads, [&](int y) {
        for (int x = 0; x < f.W; ++x) {
            Px4 cpx = in.at(x, y);
            out[(size_t)y * f.W + x] = cpx;
            F3 colorValue = YcocgToRgb(RgbToYcocg(rgb(cpx)));
            int maskValue = cpx.w;
            F3 nMax = f3(FLT_MIN), nMin = f3(FLT_MAX), nMax2 = f3(FLT_MIN), nMin2 = f3(FLT_MAX);
            for (int j = 0; j < 9; ++j) {
                int xo = j % 3, yo = j / 3;
                Px4 q = in.at(x + xo - 1, y + yo - 1);
                F3 color = RgbToYcocg(rgb(q));
                if ((int)q.w == maskValue) {
                    nMax = max3f(nMax, color); nMin = min3f(nMin, color);
                    if (abs(xo - 1) + abs(yo - 1) <= 1) { nMax2 = max3f(nMax2, color); nMin2 = min3f(nMin2, color); }
                }
            }
            nMax = (nMax + nMax2) / 2.0f; nMin = (nMin + nMin2) / 2.0f;
            const h16* mv = f.motion + 2 * ((size_t)y * f.W + x);
            F2 motionVec = F2{h2f(mv[0]), h2f(mv[1])} - 0.5f;
            F2 uv = (F2{(float)x, (float)y} + 0.5f) * F2{1.0f / (float)f.W, 1.0f / (float)f.H};
            F2 huv = uv + motionVec;
            if (huv.x < 0 || huv.y < 0 || huv.x > 1.0 || huv.y > 1.0) continue;
            F3 ch = sample_history(hist, huv);
            F3 hy = RgbToYcocg(ch);
            hy = F3{clampf(hy.x, nMin.x, nMax.x), clampf(hy.y, nMin.y, nMax.y), clampf(hy.z, nMin.z, nMax.z)};
            ch = YcocgToRgb(hy);
            float lumaMin = nMin.x, lumaMax = nMax.x, lumaCurrent = RgbToYcocg(colorValue).x;
            float discard = 0;
            int hx = (int)floorf(huv.x * hist.W), hyy = (int)floorf(huv.y * hist.H);
            for (int i = 0; i < 4; ++i) discard += (maskValue != (int)hist.at(hx + i % 2, hyy + i / 2).w);
            discard /= 4.0f;
            if (discard == 1.0f) continue;
            ch = ch * (1.0f - discard) + colorValue * discard;
            float lumaHistory = RgbToYcocg(ch).x;
            F3 o = temporal_blend(colorValue, ch, lumaHistory, lumaMin, lumaMax, lumaCurrent, 3.0f / 4.0f);
            if (nan3(o)) o = f3(0.0f);
            out[(size_t)y * f.W + x] = pack(o, (h16)maskValue);
        }
    });
}

}  // namespace orc

using namespace orc;

extern "C" {

// In/out buffers are the reference's surface formats, linear: color/accum/history = half3+u16 (8 B/px), normal/albedo =
// half4, depth = half, motion = half2.  color is filtered "in place" from the caller's point of view (final result is
// written back to `color`); accum / historyColor / historyDepth are read as the previous frame's and overwritten.
// passes: [enableTemporalDenoising, enableLocalSpatialFilter, enableWideSpatialFilter, enableTemporalDenoising2]
// Dynamic resolution (src/kernel.cu:83-84,266): accum / historyColor are the previous frame's images at THEIR sizes
// (accumW x accumH, histW x histH = historyDim of src/denoising.cu:62,178); the images of this frame are written, compact at W x H,
// to accumOut / historyColorOut / historyDepthOut (may alias the inputs only when the sizes are equal).
int orc_denoise_hist(int W, int H, int accumW, int accumH, int histW, int histH, int frameNum, const float* params11, const int* passes4,
                     void* color, const void* normal, const void* depth, const void* motion, const void* albedo,
                     const void* accum, const void* historyColor, void* accumOut, void* historyColorOut, void* historyDepthOut,
                     void* noise8, void* noise16, int threads) {
    DenoiseParams p{params11[0], params11[1], params11[2], params11[3], params11[4], params11[5], params11[6], params11[7], params11[8], params11[9], params11[10]};
    Frame f{W, H, (const Px4*)normal, (const h16*)depth, (const h16*)motion, (const Px4*)albedo};
    const size_t n = (size_t)W * H;
    if (threads < 1) threads = 1;
    std::vector<Px4> a((Px4*)color, (Px4*)color + n), b(n);
    Px4* cur = a.data(); Px4* nxt = b.data();
    auto swap = [&]() { Px4* t = cur; cur = nxt; nxt = t; };
    Img accumImg{accumW, accumH, (const Px4*)accum}, histImg{histW, histH, (const Px4*)historyColor};
    if (passes4[0] && frameNum != 1) { temporal_filter(f, Img{W, H, cur}, accumImg, nxt, p, threads); swap(); }
    if (passes4[1]) {
        noise_levels(f, Img{W, H, cur}, (h16*)noise8, (h16*)noise16);
        spatial7x7(f, Img{W, H, cur}, nxt, (const h16*)noise16, frameNum, p, threads); swap();
    }
    if (passes4[0]) std::memcpy(accumOut, cur, n * sizeof(Px4));                       // CopyToHistoryColorBuffer
    if (passes4[2]) {
        noise_levels(f, Img{W, H, cur}, (h16*)noise8, (h16*)noise16);
        // passes4[2] == 2: the 5-iteration a-trous schedule of BASELINE config 4 (strides 1, 2, 4, 8, 16: the same kernel template,
        // src/temporalDenoising.cuh:495-608, instantiated with the classic power-of-two strides instead of the shipped 3, 6, 12)
        if (passes4[2] == 2) { for (int S : {1, 2, 4, 8, 16}) { spatial5x5(f, Img{W, H, cur}, nxt, (const h16*)noise16, S, p, threads); swap(); } }
        else { for (int S : {3, 6, 12}) { spatial5x5(f, Img{W, H, cur}, nxt, (const h16*)noise16, S, p, threads); swap(); } }
    }
    apply_albedo(f, Img{W, H, cur}, nxt, threads); swap();
    if (passes4[3]) {
        if (frameNum != 1) { temporal_filter2(f, Img{W, H, cur}, histImg, nxt, threads); swap(); }
        std::memcpy(historyColorOut, cur, n * sizeof(Px4));                             // CopyToHistoryColorDepthBuffer
        std::memcpy(historyDepthOut, depth, n * sizeof(h16));
    }
    std::memcpy(color, cur, n * sizeof(Px4));
    return 0;
}

int orc_denoise(int W, int H, int frameNum, const float* params11, const int* passes4,
                void* color, const void* normal, const void* depth, const void* motion, const void* albedo,
                void* accum, void* historyColor, void* historyDepth, void* noise8, void* noise16, int threads) {
    return orc_denoise_hist(W, H, W, H, W, H, frameNum, params11, passes4, color, normal, depth, motion, albedo, accum, historyColor,
                            accum, historyColor, historyDepth, noise8, noise16, threads);
}

}  // extern "C"
