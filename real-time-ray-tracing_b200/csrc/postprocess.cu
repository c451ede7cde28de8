// §8(f) N2 — post-processing chain for sm_100a (rtrt_postprocess): what RayTracer::draw runs right after the denoiser.
//
// Restates RayTracer::PostProcessing (src/postprocessing.cu:5-160) with the default pass set of RenderPassSettings
// (src/settingParams.h:92-99: down-scale passes, histogram, auto exposure, sharpening, tone mapping on; bloom and lens flare
// off) plus the CopyToOutput launch of draw() (src/kernel.cu:26-59, :369-375):
//   DownScale4 x3 (src/postprocessing.cuh:142-174)  -> 1/64-resolution colour
//   Histogram2 (:24-39) + AutoExposure (:43-138)    -> exposure[4] (EV, adapted luminance, adapted bright luminance, bright luminance)
//   BicubicScale (:785-801, Catmull-Rom, src/sampler.cuh:445-496) render size -> output size
//   SharpeningFilter (:726-783, FidelityFX-CAS style) , ToneMappingReinhardExtended (:542-564) , CopyToOutput (blue-noise dither -> RGBA8)
// with the same formulas and the same half-precision store points, but
//   * 3 launches instead of 9: one kernel reduces 64x64 pixel tiles through the three down-scale levels in shared memory (the
//     1/4 and 1/16 images are only inputs of the disabled bloom pass and are not written), one CTA does histogram + exposure, and
//     one kernel does scale + sharpen + tone map + dither out of an 18x18 shared-memory tile, so the output-size half4 image the
//     reference writes and re-reads four times (32 B/px each way) never touches HBM;
//   * the sharpening pass reads un-sharpened neighbours (the reference filters in place while neighbouring threads read).
// Tolerance against the reference's kernels is stated in tests/test_postprocess_gpu.py.
#include "context.h"
#include "rtrt_vec.cuh"
#include <cuda_fp16.h>

namespace rtrt {

struct PpParams {
    int W, H;                 // render size
    int oW, oH;               // output (screen) size
    int w64, h64;             // 1/64-resolution size: ceil(ceil(ceil(W/4)/4)/4)
    float deltaTimeMs, gain, maxWhite, gamma, fixedExposure;
    int autoExposure, sharpening, toneMapping, postProcess;
    int frameNum;
    const ushort4* __restrict__ color;    // denoised frame, half3 + mask
    ushort4* color64;                     // half4, w64 x h64
    uint32_t* histogram;                  // 64 bins
    float* exposure;                      // 4 floats, persistent across frames
    uchar4* out;                          // oW x oH RGBA8
    const unsigned char* __restrict__ sobol; const unsigned char* __restrict__ scramble; const unsigned char* __restrict__ rank;
};

VD float pp_h2f(unsigned short b) { return __half2float(__ushort_as_half(b)); }
VD unsigned short pp_f2h(float f) { return __half_as_ushort(__float2half_rn(f)); }
VD float pp_rh(float f) { return __half2float(__float2half_rn(f)); }                    // one Store2DHalf4 / Load2DHalf4 round trip
VD Float3 pp_rh3(Float3 c) { return Float3{pp_rh(c.x), pp_rh(c.y), pp_rh(c.z)}; }
VD Float3 min3f(Float3 a, Float3 b) { return Float3{sel_min(a.x, b.x), sel_min(a.y, b.y), sel_min(a.z, b.z)}; }
VD Float3 pp_load(const ushort4* __restrict__ img, int W, int H, int x, int y) {         // surf2Dread, cudaBoundaryModeClamp
    x = clampi(x, 0, W - 1); y = clampi(y, 0, H - 1);
    const ushort4 p = __ldg(img + (size_t)y * W + x);
    return Float3{pp_h2f(p.x), pp_h2f(p.y), pp_h2f(p.z)};
}

// -------------------------------------------------------------------------------------------------------------------
// Down-scale: one CTA = one 64x64 render-pixel tile = 16x16 texels of level 1, 4x4 of level 2, 1 of level 3.
// Every level is DownScale4's 8x8-block arithmetic: v/16, 2x2 sums (left-assoc, x then y neighbours), 2x2 sums of those,
// stored as half.  Reads beyond an image clamp to its last row/column like the surface reads.
// -------------------------------------------------------------------------------------------------------------------
VD Float3 ds_sum4(Float3 a, Float3 b, Float3 c, Float3 d) { return ((a + b) + c) + d; }

__global__ void __launch_bounds__(256) k_pp_downscale64(const PpParams P) {
    __shared__ Float3 s1[16][16];          // level-1 texels of this tile (half-rounded)
    __shared__ Float3 s2[4][4];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int w1 = (P.W + 3) / 4, h1 = (P.H + 3) / 4, w2 = (w1 + 3) / 4, h2 = (h1 + 3) / 4;
    {   // level 1: thread (tx,ty) reduces render pixels [4*X1, 4*X1+4) x [4*Y1, 4*Y1+4)
        const int X1 = blockIdx.x * 16 + tx, Y1 = blockIdx.y * 16 + ty;
        Float3 q[2][2];
#pragma unroll
        for (int by = 0; by < 2; ++by)
#pragma unroll
            for (int bx = 0; bx < 2; ++bx) {
                const int x = X1 * 4 + bx * 2, y = Y1 * 4 + by * 2;
                q[bx][by] = ds_sum4(pp_load(P.color, P.W, P.H, x, y) / 16.0f, pp_load(P.color, P.W, P.H, x + 1, y) / 16.0f,
                                    pp_load(P.color, P.W, P.H, x, y + 1) / 16.0f, pp_load(P.color, P.W, P.H, x + 1, y + 1) / 16.0f);
            }
        s1[tx][ty] = pp_rh3(ds_sum4(q[0][0], q[1][0], q[0][1], q[1][1]));
    }
    __syncthreads();
    // level-1 image has w1 x h1 texels; texels beyond it clamp to the last valid one (the level-2 launch reads ColorBuffer4 clamped)
    auto l1 = [&](int X, int Y) {           // X,Y = level-1 coordinates
        X = min(X, w1 - 1) - blockIdx.x * 16; Y = min(Y, h1 - 1) - blockIdx.y * 16;
        return s1[max(X, 0)][max(Y, 0)];
    };
    if (threadIdx.x < 16) {
        const int ux = threadIdx.x & 3, uy = threadIdx.x >> 2;
        const int X2 = blockIdx.x * 4 + ux, Y2 = blockIdx.y * 4 + uy;
        Float3 q[2][2];
#pragma unroll
        for (int by = 0; by < 2; ++by)
#pragma unroll
            for (int bx = 0; bx < 2; ++bx) {
                const int x = X2 * 4 + bx * 2, y = Y2 * 4 + by * 2;
                q[bx][by] = ds_sum4(l1(x, y) / 16.0f, l1(x + 1, y) / 16.0f, l1(x, y + 1) / 16.0f, l1(x + 1, y + 1) / 16.0f);
            }
        s2[ux][uy] = pp_rh3(ds_sum4(q[0][0], q[1][0], q[0][1], q[1][1]));
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        auto l2 = [&](int X, int Y) {
            X = min(X, w2 - 1) - blockIdx.x * 4; Y = min(Y, h2 - 1) - blockIdx.y * 4;
            return s2[max(X, 0)][max(Y, 0)];
        };
        const int X3 = blockIdx.x, Y3 = blockIdx.y;
        Float3 q[2][2];
#pragma unroll
        for (int by = 0; by < 2; ++by)
#pragma unroll
            for (int bx = 0; bx < 2; ++bx) {
                const int x = X3 * 4 + bx * 2, y = Y3 * 4 + by * 2;
                q[bx][by] = ds_sum4(l2(x, y) / 16.0f, l2(x + 1, y) / 16.0f, l2(x, y + 1) / 16.0f, l2(x + 1, y + 1) / 16.0f);
            }
        const Float3 r = ds_sum4(q[0][0], q[1][0], q[0][1], q[1][1]);
        if (X3 < P.w64 && Y3 < P.h64) P.color64[(size_t)Y3 * P.w64 + X3] = make_ushort4(pp_f2h(r.x), pp_f2h(r.y), pp_f2h(r.z), pp_f2h(0.0f));
    }
}

// -------------------------------------------------------------------------------------------------------------------
// Histogram2 (one block of min(w64,32) x min(h64,32) threads over the 1/64 image) + AutoExposure (one thread).
// -------------------------------------------------------------------------------------------------------------------
VD float bin_to_lum(int i) { return exp2f(((float)i / (63 * 0.99999) - 0.75) / 0.1); }   // double constants like the reference

__global__ void __launch_bounds__(1024) k_pp_exposure(const PpParams P) {
    __shared__ uint32_t hist[64];
    const int tid = threadIdx.y * blockDim.x + threadIdx.x, nt = blockDim.x * blockDim.y;
    for (int i = tid; i < 64; i += nt) hist[i] = 0u;
    __syncthreads();
    if ((int)threadIdx.x < P.w64 && (int)threadIdx.y < P.h64) {
        const ushort4 p = P.color64[(size_t)threadIdx.y * P.w64 + threadIdx.x];
        const Float3 c{pp_h2f(p.x), pp_h2f(p.y), pp_h2f(p.z)};
        const float luminance = dot(c, Float3{0.3f, 0.6f, 0.1f});                    // compensated dot (src/linearMath.h:484)
        const float logLuminance = (float)(log2f(luminance) * 0.1 + 0.75);
        const uint32_t bucket = (uint32_t)rintf((float)(clampf(logLuminance, 0.0f, 1.0f) * 63 * 0.99999));
        atomicAdd(&hist[bucket], 1u);
    }
    __syncthreads();
    for (int i = tid; i < 64; i += nt) P.histogram[i] = hist[i];
    if (tid != 0) return;
    if (!P.autoExposure) {                                                           // src/postprocessing.cu:52-60
        P.exposure[0] = P.fixedExposure; P.exposure[1] = 1.0f; P.exposure[2] = 1.0f; P.exposure[3] = 1.0f;
        return;
    }
    const float area = (float)(P.w64 * P.h64);
    const float darkThreshold = 0.4f, brightThreshold = 0.9f;
    float lumiSum = 0, lumiSumArea = 0, accuHistArea = 0, brightLum = 0;
    int i = 0;
    for (; i < 64; ++i) {
        const float fHist = (float)hist[i] / area, lum = bin_to_lum(i);
        accuHistArea += fHist;
        const float dark = accuHistArea - darkThreshold;
        if (dark > 0) { lumiSumArea += dark; lumiSum += dark * lum; break; }
    }
    for (; i < 64; ++i) {
        const float fHist = (float)hist[i] / area, lum = bin_to_lum(i);
        accuHistArea += fHist;
        const float bright = accuHistArea - brightThreshold;
        if (bright > 0) {
            const float partial = brightThreshold - (accuHistArea - fHist);
            lumiSumArea += partial; lumiSum += partial * lum; brightLum = lum;
            break;
        } else { lumiSumArea += fHist; lumiSum += fHist * lum; }
    }
    float aveLum = clampf(lumiSum / lumiSumArea, 0.1f, 100.0f);
    float lumTemp = P.exposure[1], lumBrightTemp = P.exposure[2];
    lumTemp = lumTemp + (aveLum - lumTemp) * (1.0f - expf(-P.deltaTimeMs * 0.001f));
    lumBrightTemp = lumBrightTemp + (brightLum - lumBrightTemp) * (1.0f - expf(-P.deltaTimeMs * 0.001f));
    const float EC = 1.03f - 2.0f / (log10f(lumTemp + 1.0f) + 2.0f);
    P.exposure[0] = P.gain * EC / lumTemp; P.exposure[1] = lumTemp; P.exposure[2] = lumBrightTemp; P.exposure[3] = brightLum;
}

// -------------------------------------------------------------------------------------------------------------------
// Scale + sharpen + tone map + dither -> RGBA8.  16x16 output pixels per CTA, 18x18 staged tile of scaled colours.
// -------------------------------------------------------------------------------------------------------------------
VD Float3 catmull_rom(const PpParams& P, int ox, int oy) {                  // BicubicScale + SampleBicubicCatmullRom
    const Float2 uv{(float)ox / P.oW, (float)oy / P.oH};
    const Float2 UV = uv * Float2{(float)P.W, (float)P.H};
    const Float2 tc{floorf(UV.x - 0.5f) + 0.5f, floorf(UV.y - 0.5f) + 0.5f};
    const Float2 f = UV - tc, f2 = f * f, fc = f2 * f;
    const Float2 w0 = f2 - (fc + f) * 0.5f;
    const Float2 w1 = fc * 1.5f - f2 * 2.5f + 1.0f;
    const Float2 w3 = (fc - f2) * 0.5f;
    const Float2 w2 = ((Float2{1.0f, 1.0f} - w0) - w1) - w3;
    const int x1 = (int)floorf(UV.x - 0.5f), y1 = (int)floorf(UV.y - 0.5f);
    const float wx[4] = {w0.x, w1.x, w2.x, w3.x}, wy[4] = {w0.y, w1.y, w2.y, w3.y};
    Float3 out = f3(0.0f); float sum = 0.0f;
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float w = wx[i] * wy[j];
            sum += w;
            out = out + pp_load(P.color, P.W, P.H, x1 - 1 + i, y1 - 1 + j) * w;
        }
    return out / sum;
}

VD float pp_blue(const PpParams& P, int px, int py, int sampleIdx, int dim) {         // src/blueNoiseRandGen.h:119-146
    px &= 127; py &= 127; sampleIdx &= 255;
    const int ranked = sampleIdx ^ __ldg(P.rank + dim + (px + py * 128) * 8);
    int v = __ldg(P.sobol + dim + ranked * 256);
    v ^= __ldg(P.scramble + (dim % 8) + (px + py * 128) * 8);
    return (v + 0.5f) / 256.0f;
}

__global__ void __launch_bounds__(256) k_pp_output(const PpParams P) {
    constexpr int T = 18;
    __shared__ Float3 sC[T * T];
    const int tid = threadIdx.y * 16 + threadIdx.x;
    const int x0 = blockIdx.x * 16 - 1, y0 = blockIdx.y * 16 - 1;
    for (int i = tid; i < T * T; i += 256) {
        const int gx = clampi(x0 + i % T, 0, P.oW - 1), gy = clampi(y0 + i / T, 0, P.oH - 1);   // SharpeningFilter's clamped neighbour reads
        sC[i] = pp_rh3(catmull_rom(P, gx, gy));                                                   // Store2DHalf4 of BicubicScale
    }
    __syncthreads();
    const int x = blockIdx.x * 16 + threadIdx.x, y = blockIdx.y * 16 + threadIdx.y;
    if (x >= P.oW || y >= P.oH) return;
    auto C = [&](int i, int j) { return sC[(threadIdx.y + 1 + j) * T + threadIdx.x + 1 + i]; };   // i,j in -1..1 = color[i+1][j+1]
    Float3 color = C(0, 0);
    if (P.postProcess && P.sharpening) {
        Float3 t1 = max3f(max3f(C(0, 0), C(-1, 0)), C(1, 0));
        Float3 t2 = max3f(max3f(t1, C(0, -1)), C(0, 1));
        Float3 t3 = max3f(max3f(t2, C(-1, -1)), C(-1, 1));
        Float3 t4 = max3f(max3f(t3, C(1, -1)), C(1, 1));
        const Float3 softmax = t2 + t4;
        t1 = min3f(min3f(C(0, 0), C(-1, 0)), C(1, 0));
        t2 = min3f(min3f(t1, C(0, -1)), C(0, 1));
        t3 = min3f(min3f(t2, C(-1, -1)), C(-1, 1));
        t4 = min3f(min3f(t3, C(1, -1)), C(1, 1));
        const Float3 softmin = t2 + t4;
        const Float3 m = min3f(softmin, f3(2.0f) - softmax) / softmax;
        Float3 amp{clampf(m.x, 0.0f, 1.0f), clampf(m.y, 0.0f, 1.0f), clampf(m.z, 0.0f, 1.0f)};
        amp = Float3{1.0f / sqrtf(amp.x), 1.0f / sqrtf(amp.y), 1.0f / sqrtf(amp.z)};
        const float peak = 8.0f - 3.0f * 1.0f;
        const Float3 w = (f3(0.0f) - f3(1.0f)) / (amp * peak);
        Float3 o = (((C(-1, 0) + C(1, 0)) + C(0, -1)) + C(0, 1)) * w + C(0, 0);
        o = o / (f3(1.0f) + w * 4.0f);
        color = pp_rh3(o);
    }
    if (P.postProcess && P.toneMapping) {                                      // ToneMappingReinhardExtended
        color = color * P.exposure[0];
        const float lOld = dot(color, Float3{0.2126f, 0.7152f, 0.0722f});
        const float numerator = lOld * (1.0f + (lOld / (P.maxWhite * P.maxWhite)));
        const float lNew = numerator / (1.0f + lOld);
        color = color * (lNew / dot(color, Float3{0.2126f, 0.7152f, 0.0722f}));
        const float ig = 1.0f / P.gamma;
        color = Float3{clampf(powf(color.x, ig), 0.0f, 1.0f), clampf(powf(color.y, ig), 0.0f, 1.0f), clampf(powf(color.z, ig), 0.0f, 1.0f)};
        color = pp_rh3(color);
    }
    // CopyToOutput (src/kernel.cu:26-59): jitter = blueNoise/256 - 1/512 with INTEGER 1/512 == 0, as written in the reference
    const Float3 bn{pp_blue(P, x, y, P.frameNum, 0), pp_blue(P, x, y, P.frameNum, 1), pp_blue(P, x, y, P.frameNum, 2)};
    const Float3 jitter = bn / 256.0f;
    const float hi = 1.0f - 1.1920928955078125e-07f;
    const Float3 s{clampf(color.x + jitter.x, 0.0f, hi), clampf(color.y + jitter.y, 0.0f, hi), clampf(color.z + jitter.z, 0.0f, hi)};
    P.out[(size_t)y * P.oW + x] = make_uchar4((unsigned char)(s.x * 256), (unsigned char)(s.y * 256), (unsigned char)(s.z * 256), 1);
}

// -------------------------------------------------------------------------------------------------------------------
int launch_postprocess(Context& c, float deltaTimeMs) {
    if (!c.dColor || c.gbufW != c.width || c.gbufH != c.height) return c.fail(RTRT_ERR_STATE, "rtrt_postprocess: no frame to process");
    const int oW = c.outW > 0 ? c.outW : c.width, oH = c.outH > 0 ? c.outH : c.height;
    const int w1 = (c.width + 3) / 4, h1 = (c.height + 3) / 4, w2 = (w1 + 3) / 4, h2 = (h1 + 3) / 4, w3 = (w2 + 3) / 4, h3 = (h2 + 3) / 4;
    if (!c.dOutRgba || c.ppOutW != oW || c.ppOutH != oH) {
        if (c.dOutRgba) cudaFree(c.dOutRgba);
        c.dOutRgba = nullptr;
        RTRT_CUDA(c, cudaMalloc(&c.dOutRgba, (size_t)oW * oH * 4));
        c.ppOutW = oW; c.ppOutH = oH;
    }
    if (!c.dColor64 || (size_t)w3 * h3 > c.ppCap64) {                                   // grow-only: the render size may change every frame
        if (c.dColor64) cudaFree(c.dColor64);
        c.dColor64 = nullptr;
        RTRT_CUDA(c, cudaMalloc(&c.dColor64, (size_t)w3 * h3 * 8));
        c.ppCap64 = (size_t)w3 * h3;
    }
    c.ppW64 = w3; c.ppH64 = h3;
    if (!c.dExposure) {
        RTRT_CUDA(c, cudaMalloc(&c.dExposure, 4 * sizeof(float) + 64 * sizeof(uint32_t)));
        const float init[4] = {1.0f, 1.0f, 1.0f, 1.0f};                                  // src/init.cu:331-333
        RTRT_CUDA(c, cudaMemcpyAsync(c.dExposure, init, sizeof(init), cudaMemcpyHostToDevice, c.stream));
        RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    }
    PpParams P{};
    P.W = c.width; P.H = c.height; P.oW = oW; P.oH = oH; P.w64 = w3; P.h64 = h3;
    P.deltaTimeMs = deltaTimeMs; P.gain = c.postParams.gain; P.maxWhite = c.postParams.maxWhite; P.gamma = c.postParams.gamma;
    P.fixedExposure = c.postParams.exposure;
    P.postProcess = c.postParams.enablePostProcess; P.autoExposure = c.postParams.enableAutoExposure;
    P.sharpening = c.postParams.enableSharpening; P.toneMapping = c.postParams.enableToneMapping;
    P.frameNum = c.frameNum;
    P.color = reinterpret_cast<const ushort4*>(c.dColor); P.color64 = reinterpret_cast<ushort4*>(c.dColor64);
    P.exposure = reinterpret_cast<float*>(c.dExposure); P.histogram = reinterpret_cast<uint32_t*>(reinterpret_cast<float*>(c.dExposure) + 4);
    P.out = reinterpret_cast<uchar4*>(c.dOutRgba);
    P.sobol = c.dBlueNoise; P.scramble = c.dBlueNoise + 65536; P.rank = c.dBlueNoise + 65536 + 131072;
    if (P.postProcess) {
        k_pp_downscale64<<<dim3(w3, h3), 256, 0, c.stream>>>(P);
        k_pp_exposure<<<1, dim3(w3 < 32 ? w3 : 32, h3 < 32 ? h3 : 32), 0, c.stream>>>(P);
        c.launches += 2;
    }
    k_pp_output<<<dim3((oW + 15) / 16, (oH + 15) / 16), dim3(16, 16), 0, c.stream>>>(P);
    c.launches += 1;
    RTRT_CUDA(c, cudaGetLastError());
    return RTRT_OK;
}

}  // namespace rtrt
