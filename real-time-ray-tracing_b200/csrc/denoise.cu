// Stage 3 — temporal + edge-aware spatial denoiser for sm_100a (rtrt_denoise).
//
// Restates RayTracer::TemporalSpatialDenoising (src/denoising.cu:5-188) and its kernels
// (src/temporalDenoising.cuh: TemporalFilter :610-893, CalculateTileNoiseLevel :33-91, TileNoiseLevel8x8to16x16
// :93-102, SpatialFilter7x7 :317-492, SpatialFilterGlobal5x5<3|6|12> :495-608, ApplyAlbedo :1127-1139,
// TemporalFilter2 :896-1111, CopyToHistory* :142-170) with the same formulas, tap sets, thresholds and the same
// half-precision store points, but
//   * PING-PONG buffers instead of in-place filtering (the reference's in-place passes race across tiles, so its
//     output is not even deterministic),
//   * 6 launches instead of 12: the two noise-level kernels are folded into the pass that produces the colours
//     they measure (tile statistics come out of shared memory, same shuffle order as WarpReduceSum), the
//     accumulation copy is a second store of the 7x7 pass, ApplyAlbedo + TemporalFilter2 + the history copies
//     are one pass.  Pass-wise traffic drops from 252 B/px to ~150 B/px.
// Buffers are linear (8-byte pixels: half3 colour + u16 mask), read through the read-only path.
#include "context.h"
#include "rtrt_vec.cuh"
#include <cuda_fp16.h>

namespace rtrt {

struct DnParams {
    int W, H, frameNum;
    int aW, aH, hW, hH;                       // historyDim of the accumulation / history colour images (src/denoising.cu:62,178)
    int by0;                                  // first 16-row block row this launch covers (row-band mode; 0 = whole frame)
    float local_n, local_d, local_m, large_n, large_d, large_m, temp_n, temp_d, temp_m, thr_local, thr_large;
    const ushort4* __restrict__ normal; const __half* __restrict__ depth; const __half2* __restrict__ motion;
    const ushort4* __restrict__ albedo;
    __half* noise8; __half* noise16;
    const ushort4* __restrict__ guide;       // normal.xyz | depth, packed once per frame by k_dn_pack_guide (8 B/px)
};

__constant__ float cG3[9] = {0.0578968, 0.0921378, 0.0584323, 0.0921378, 0.146629, 0.09299, 0.0584322, 0.0929898, 0.0589727};
__constant__ float cG5[25] = {0.00360466, 0.0144464, 0.0229902, 0.01458, 0.0036719, 0.0144464, 0.0578968, 0.0921378, 0.0584323, 0.0147159,
                              0.0229902, 0.0921378, 0.146629, 0.09299, 0.023419, 0.01458, 0.0584322, 0.0929898, 0.0589727, 0.014852,
                              0.00367191, 0.0147158, 0.0234191, 0.0148519, 0.0037404};
// the same tables as compile-time constants: immediates in the fully unrolled tap loops (the __constant__ copies serve the rolled loops)
__device__ constexpr float kG5[25] = {0.00360466, 0.0144464, 0.0229902, 0.01458, 0.0036719, 0.0144464, 0.0578968, 0.0921378, 0.0584323, 0.0147159,
                                      0.0229902, 0.0921378, 0.146629, 0.09299, 0.023419, 0.01458, 0.0584322, 0.0929898, 0.0589727, 0.014852,
                                      0.00367191, 0.0147158, 0.0234191, 0.0148519, 0.0037404};
__device__ constexpr float kG3[9] = {0.0578968, 0.0921378, 0.0584323, 0.0921378, 0.146629, 0.09299, 0.0584322, 0.0929898, 0.0589727};
__device__ constexpr float kLg2G3[9] = {-4.11037254f, -3.440063f, -4.09709024f, -3.440063f, -2.76975751f, -3.42678046f, -4.09709263f, -3.42678356f, -4.0838089f};
__device__ constexpr float kG7[49] = {3.47404e-05, 0.000353875, 0.00141822, 0.00225698, 0.00143134, 0.000360475, 3.57221e-05,
                                      0.000353875, 0.00360466, 0.0144464, 0.0229902, 0.01458, 0.0036719, 0.000363875,
                                      0.00141822, 0.0144464, 0.0578968, 0.0921378, 0.0584323, 0.0147159, 0.0014583,
                                      0.00225698, 0.0229902, 0.0921378, 0.146629, 0.09299, 0.023419, 0.00232076,
                                      0.00143134, 0.01458, 0.0584322, 0.0929898, 0.0589727, 0.014852, 0.00147179,
                                      0.000360475, 0.00367191, 0.0147158, 0.0234191, 0.0148519, 0.0037404, 0.000370662,
                                      3.57221e-05, 0.000363875, 0.0014583, 0.00232075, 0.00147179, 0.000370662, 3.67315e-05};
// log2 of the tables (RTRT_DN_FMA: the Gaussian rides in the exponent)
__device__ constexpr float kLg2G7[49] = {-14.8130264f, -11.4644728f, -9.4617033f, -8.79139042f, -9.44841766f, -11.4378128f, -14.7728233f,
                                          -11.4644728f, -8.11592102f, -6.11314631f, -5.44283724f, -6.09986544f, -8.08925724f, -11.4242697f,
                                          -9.4617033f, -6.11314631f, -4.11037254f, -3.440063f, -4.09709024f, -6.08648062f, -9.42149639f,
                                          -8.79139042f, -5.44283724f, -3.440063f, -2.76975751f, -3.42678046f, -5.4161768f, -8.75118732f,
                                          -9.44841766f, -6.09986544f, -4.09709263f, -3.42678356f, -4.0838089f, -6.0731988f, -9.40821266f,
                                          -11.4378128f, -8.08925343f, -6.08649015f, -5.4161706f, -6.07320881f, -8.06259155f, -11.3976078f,
                                          -14.7728233f, -11.4242697f, -9.42149639f, -8.75119305f, -9.40821266f, -11.3976078f, -14.7326231f};
__device__ constexpr float kLg2G5[25] = {-8.11592102f, -6.11314631f, -5.44283724f, -6.09986544f, -8.08925724f,
                                          -6.11314631f, -4.11037254f, -3.440063f, -4.09709024f, -6.08648062f,
                                          -5.44283724f, -3.440063f, -2.76975751f, -3.42678046f, -5.4161768f,
                                          -6.09986544f, -4.09709263f, -3.42678356f, -4.0838089f, -6.0731988f,
                                          -8.08925343f, -6.08649015f, -5.4161706f, -6.07320881f, -8.06259155f};
__constant__ float cG7[49] = {3.47404e-05, 0.000353875, 0.00141822, 0.00225698, 0.00143134, 0.000360475, 3.57221e-05,
                              0.000353875, 0.00360466, 0.0144464, 0.0229902, 0.01458, 0.0036719, 0.000363875,
                              0.00141822, 0.0144464, 0.0578968, 0.0921378, 0.0584323, 0.0147159, 0.0014583,
                              0.00225698, 0.0229902, 0.0921378, 0.146629, 0.09299, 0.023419, 0.00232076,
                              0.00143134, 0.01458, 0.0584322, 0.0929898, 0.0589727, 0.014852, 0.00147179,
                              0.000360475, 0.00367191, 0.0147158, 0.0234191, 0.0148519, 0.0037404, 0.000370662,
                              3.57221e-05, 0.000363875, 0.0014583, 0.00232075, 0.00147179, 0.000370662, 3.67315e-05};

VD float h2f(unsigned short b) { return __half2float(__ushort_as_half(b)); }
VD unsigned short f2h(float f) { return __half_as_ushort(__float2half_rn(f)); }
VD Float3 px_rgb(ushort4 p) { return Float3{h2f(p.x), h2f(p.y), h2f(p.z)}; }
VD ushort4 px_pack(Float3 c, unsigned short w) { return make_ushort4(f2h(c.x), f2h(c.y), f2h(c.z), w); }
VD Float3 min3f(Float3 a, Float3 b) { return Float3{sel_min(a.x, b.x), sel_min(a.y, b.y), sel_min(a.z, b.z)}; }
VD Float3 rgb_to_ycocg(Float3 c) { const float t1 = c.x + c.z, t2 = c.y * 2.0f; return Float3{t1 + t2, (c.x - c.z) * 2.0f, t2 - t1}; }
VD Float3 ycocg_to_rgb(Float3 y) { const float t = y.x - y.z; return Float3{t + y.y, y.x + y.z, t - y.y} * 0.25f; }
VD ushort4 ld_px(const ushort4* __restrict__ img, int W, int H, int x, int y) {      // surface clamp addressing
    x = clampi(x, 0, W - 1); y = clampi(y, 0, H - 1);
    return __ldg(img + (size_t)y * W + x);
}

// SampleBicubicSmoothStep<Load2DFuncHalf3Ushort1>, default boundary (src/sampler.cuh:498-541)
VD Float3 sample_history(const ushort4* __restrict__ img, int W, int H, Float2 uv) {
    const BicubicTaps t = bicubic_taps(uv, W, H);
    Float3 out = f3(0.0f);
#pragma unroll
    for (int i = 0; i < 4; ++i) out = out + px_rgb(ld_px(img, W, H, t.x0 + (i & 1), t.y0 + (i >> 1))) * t.w[i];
    return out / t.sum;
}

// blend tail shared by TemporalFilter (:861-889) and TemporalFilter2 (:1079-1107)
VD Float3 temporal_blend(Float3 colorValue, Float3 hist, float lumaHistory, float lumaMin, float lumaMax, float lumaCurrent, float base) {
    float blend = base;
    blend *= 0.2f + 0.8f * clampf(0.5f * sel_min(fabsf(lumaHistory - lumaMin), fabsf(lumaHistory - lumaMax)) / sel_max(sel_max(lumaHistory, lumaCurrent), 1e-4f));
    float wA = blend * sel_max(0.0001f, 1.0f / (lumaCurrent + 4.0f));
    float wB = (1.0f - blend) * sel_max(0.0001f, 1.0f / (lumaHistory + 4.0f));
    const float s = wA + wB;
    const float ws = 1.0f / ((fabsf(s) > 1e-20f) ? s : copysignf(1e-20f, s));            // SafeDivide
    wA *= ws; wB *= ws;
    return colorValue * wA + hist * wB;
}

// ---- tile noise statistics out of shared memory: CalculateTileNoiseLevel (:33-91) + 8x8 -> 16x16 (:93-102) ----
// sOut = the 16x16 block's output colours (half-rounded, as stored); sDepth = its depths.  Warps 0..3 take one 8x8
// tile each with the reference's lane mapping and shfl_down tree; thread 0 then averages to the 16x16 level.
__device__ __forceinline__ void block_noise_levels(const DnParams& P, const ushort4* sOut, const unsigned short* sDepth, float* sN8, int tid) {
    const int warp = tid >> 5, lane = tid & 31;
    const int gw = (P.W + 7) / 8, gh = (P.H + 7) / 8;
    if (warp < 4) {
        const int t8x = blockIdx.x * 2 + (warp & 1), t8y = (blockIdx.y + P.by0) * 2 + (warp >> 1);
        const int tx = lane & 7, ty = lane >> 3;
        float l1, l2, b1, b2;
        {
            // pixel coordinates are clamped to the image like the surface reads, then made block-local
            const int gx = min(t8x * 8 + tx, P.W - 1) - blockIdx.x * 16;
            const int gy1 = min(t8y * 8 + 2 * ty, P.H - 1) - (blockIdx.y + P.by0) * 16, gy2 = min(t8y * 8 + 2 * ty + 1, P.H - 1) - (blockIdx.y + P.by0) * 16;
            const bool in = gx >= 0 && gx < 16 && gy1 >= 0 && gy1 < 16 && gy2 >= 0 && gy2 < 16;   // false only for tiles fully outside the image
            const ushort4 p1 = in ? sOut[gy1 * 16 + gx] : make_ushort4(0, 0, 0, 0), p2 = in ? sOut[gy2 * 16 + gx] : make_ushort4(0, 0, 0, 0);
            l1 = getmax3(px_rgb(p1)); l2 = getmax3(px_rgb(p2));
            b1 = (in && h2f(sDepth[gy1 * 16 + gx]) >= kRayMax) ? 1.0f : 0.0f;
            b2 = (in && h2f(sDepth[gy2 * 16 + gx]) >= kRayMax) ? 1.0f : 0.0f;
        }
        float l12 = l1 * l1, l22 = l2 * l2;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            b1 += __shfl_down_sync(0xffffffffu, b1, o); b2 += __shfl_down_sync(0xffffffffu, b2, o);
            l1 += __shfl_down_sync(0xffffffffu, l1, o); l12 += __shfl_down_sync(0xffffffffu, l12, o);
            l2 += __shfl_down_sync(0xffffffffu, l2, o); l22 += __shfl_down_sync(0xffffffffu, l22, o);
        }
        if (lane == 0) {
            const float notSky = 1.0f - (b1 + b2) / 64.0f;
            const float ave = (l1 + l2) / 64.0f, aveSq = ave * ave, sqAve = (l12 + l22) / 64.0f;
            const float var = sel_max(1e-20f, sqAve - aveSq);
            float noise = var / sel_max(aveSq, 1e-20f);
            noise *= notSky;
            const __half hn = __float2half_rn(noise);
            sN8[warp] = __half2float(hn);
            if (t8x < gw && t8y < gh) P.noise8[(size_t)t8y * gw + t8x] = hn;
        }
    }
    __syncthreads();
    if (tid == 0) {
        // tiles beyond the 8x8 grid read the clamped (= last valid) tile, like the surface read in :97-100
        const bool xo = blockIdx.x * 2 + 1 >= gw, yo = (blockIdx.y + P.by0) * 2 + 1 >= gh;
        const float v1 = sN8[0], v2 = sN8[xo ? 0 : 1], v3 = sN8[yo ? 0 : 2], v4 = sN8[(yo ? 0 : 2) + (xo ? 0 : 1)];
        P.noise16[(size_t)(blockIdx.y + P.by0) * ((P.W + 15) / 16) + blockIdx.x] = __float2half_rn((v1 + v2 + v3 + v4) / 4);
    }
}

// edge-stopping weight shared by the three bilateral filters (src/temporalDenoising.cuh:456-460, 565-569, 742-746).
// powf(x, sigma) and expf(x) are evaluated as 2^(sigma*log2 x) and 2^(x*log2 e) on the SFU (MUFU.LG2 / MUFU.EX2):
// three instructions instead of ~150 per tap, for 9 + 24 + 75 taps per pixel.  Relative error <= ~sigma * 2^-21
// (1e-4 at sigma = 100) against the denoiser's stated 2e-2 rel-RMSE tolerance; tests/test_denoise_gpu.py holds.
// plain 3-term dot for the edge weights (the reference's compensated dot, src/linearMath.h:484, costs ~30 instructions per tap;
// the difference is one ulp of a value that is then raised to a power with 1e-4 relative error anyway)
VD float dot_plain(Float3 a, Float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// The two factors share one exponential:  max(dot,floor)^sigN * exp(-dd^2/2) = 2^(sigN*log2(max(dot,floor)) - (log2 e / 2) dd^2)
// with dd = (d0 - d1) * (1/sigD): one MUFU.LG2, one MUFU.EX2 and no division per tap.  BilateralK holds the per-filter constants
// (1/sigD and 1/sigM are rounded once per thread; the reference divides per tap, which differs by an ulp of dd — far inside 2e-2).
struct BilateralK { float sigN, invD, invM, floorDot; };
VD BilateralK bilateral_k(float sigN, float sigD, float sigM, float floorDot) { return BilateralK{sigN, 1.0f / sigD, 1.0f / sigM, floorDot}; }
VD float fast_lg2(float x) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
VD float fast_ex2(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

VD float bilateral_weight(Float3 n0, Float3 n1, float d0, float d1, bool maskDiffers, const BilateralK& k, float g) {
    const float dn = sel_max(dot_plain(n0, n1), k.floorDot);
    const float ln = (k.sigN == 0.0f) ? 0.0f : k.sigN * fast_lg2(dn);            // x^0 = 1 also for x = 0 (log2 0 = -inf)
    const float dd = (d0 - d1) * k.invD;
    float w = fast_ex2(ln - 0.72134752044448170368f * (dd * dd));                // log2(e) / 2
    w *= maskDiffers ? k.invM : 1.0f;
    w *= g;
    return w;
}

// -------------------------------------------------------------------------------------------------------------------
// Guide pack: normal.xyz (half) + depth (half) -> one 8-byte texel, so every later tap costs one guide load, not two.
// -------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_dn_pack_guide(const ushort4* __restrict__ normal, const __half* __restrict__ depth, ushort4* __restrict__ guide, size_t first, size_t n) {
    const size_t i = first + (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= first + n) return;
    ushort4 g = __ldg(normal + i);
    g.w = __half_as_ushort(__ldg(depth + i));
    guide[i] = g;
}

// -------------------------------------------------------------------------------------------------------------------
// Tile staging through the TMA unit.  The two spatially filtered passes with a shared-memory tile (TemporalFilter: 16x16 pixels
// + 1-pixel halo, SpatialFilter7x7: + 3-pixel halo) read their colour and guide tiles with one bulk asynchronous copy
// (cp.async.bulk, global -> shared, completion on an mbarrier) per tile row and image instead of two loads, two stores and the
// clamp arithmetic per pixel and thread.  A bulk copy needs 16-byte aligned addresses and sizes and the pixels are 8 bytes, so the
// staged row starts one pixel left of the halo (an even x) and is T + 2 pixels long; the tile is addressed with that pitch (TP)
// and offset (OX = 1) on both paths.  Rows are clamped to the image per copy (surface clamp addressing in y); CTAs whose row
// range crosses the left / right image border need per-pixel clamping in x and take the per-thread path, as do gated tiles of
// SpatialFilter7x7 (they only need their own 16x16 pixels) and images with an odd width.
// -------------------------------------------------------------------------------------------------------------------
#ifndef RTRT_DN_MINB
#define RTRT_DN_MINB 6          // <= 40 registers, 6 CTAs/SM: 0.360 ms at 1080p (5: 0.362, 8: 0.360; a bound of 1 makes ptxas pick a slower schedule: 0.389)
#endif
#ifndef RTRT_DN_TMA
#define RTRT_DN_TMA 1          // 0: every tile through the per-thread path (A/B and debugging)
#endif
VD uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int T, int HALO>
struct TileStage {
    static_assert((HALO & 1) == 1 && (T & 1) == 0, "the aligned row start below assumes an odd halo and an even tile size");
    static constexpr int OX = 1, TP = T + 2;
    static constexpr uint32_t kRowBytes = TP * sizeof(ushort4);
    // lo..hi = the part of the tile (in tile coordinates) that the per-thread path must fill
    static __device__ __forceinline__ void load(const DnParams& P, const ushort4* __restrict__ in, ushort4* sC, ushort4* sG, unsigned long long* mbar,
                                                int tid, int bx, int by, int lo, int hi, bool wantTma) {
        const int x0 = bx * 16 - HALO, y0 = by * 16 - HALO, xa = x0 - OX;
        const bool tma = RTRT_DN_TMA && wantTma && (P.W & 1) == 0 && xa >= 0 && xa + TP <= P.W &&
                         ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(P.guide)) & 15) == 0;      // uniform per CTA
        if (tma) {
            if (tid == 0) {
                asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(mbar)));
                asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            }
            __syncthreads();
            if (tid == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(mbar)), "r"(2u * T * kRowBytes) : "memory");
            if (tid < 2 * T) {
                const int img = tid >= T, ly = tid - img * T;
                const int gy = clampi(y0 + ly, 0, P.H - 1);
                const ushort4* src = (img ? P.guide : in) + (size_t)gy * P.W + xa;
                ushort4* dst = (img ? sG : sC) + ly * TP;
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             :: "r"(smem_u32(dst)), "l"(src), "r"(kRowBytes), "r"(smem_u32(mbar)) : "memory");
            }
            uint32_t done = 0;
            while (!done) {
                asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                             : "=r"(done) : "r"(smem_u32(mbar)) : "memory");
            }
        } else {
            for (int i = tid; i < T * T; i += 256) {
                const int lx = i % T, ly = i / T;
                if (lx < lo || lx >= hi || ly < lo || ly >= hi) continue;
                const int gx = clampi(x0 + lx, 0, P.W - 1), gy = clampi(y0 + ly, 0, P.H - 1);
                const size_t g = (size_t)gy * P.W + gx;
                sC[ly * TP + lx + OX] = __ldg(in + g);
                sG[ly * TP + lx + OX] = __ldg(P.guide + g);
            }
            __syncthreads();
        }
    }
};

// -------------------------------------------------------------------------------------------------------------------
// Fast tap paths shared by the three bilateral filters
// -------------------------------------------------------------------------------------------------------------------
#ifndef RTRT_DN_FAST
#define RTRT_DN_FAST 1          // tiles whose taps all lie inside the image: tap addresses = row pointer + immediate, no clamps, and the per-tap
                                // NaN test of the colour moved behind the loop (a NaN tap makes the sum NaN: the pixel is then redone by the exact
                                // loop).  Same operations in the same order for every pixel: results bit-identical with RTRT_DN_FAST 0.
#endif
#ifndef RTRT_DN_FMA
#define RTRT_DN_FMA 1           // 1: the fast paths use explicit FMAs (dot, exponent, accumulation), carry log2 of the Gaussian in the exponent and divide
                                // the sums through one reciprocal: weights differ by ~1e-6 relative from RTRT_DN_FMA 0 (which is bit-identical with the
                                // exact loops), against the denoiser's 2e-2 rel-RMSE bound vs the reference kernels.  B200, 1080p: 0.307 -> 0.285 ms.
#endif

// weight of the fast paths: RTRT_DN_FMA 0 = bilateral_weight (bit-identical with the exact loops); 1 = explicit FMAs, log2 of the Gaussian added
// to the exponent, x^0 = 1 through the floor (max(dot, 1) -> log2 = 0) instead of a select
VD float floor_eff(const BilateralK& k) { return (RTRT_DN_FMA && k.sigN == 0.0f) ? 1.0f : k.floorDot; }
VD void tap_fast(ushort4 q, ushort4 gg, Float3 n0, float d0, unsigned short maskValue, const BilateralK& k, float floorEff, float g, float lg2g, Float3& sumC, float& sumW) {
    const Float3 color = px_rgb(q), n1 = px_rgb(gg);
#if RTRT_DN_FMA
    const float dn = sel_max(fmaf(n0.z, n1.z, fmaf(n0.y, n1.y, n0.x * n1.x)), floorEff);
    const float dd = (d0 - h2f(gg.w)) * k.invD;
    const float e = fmaf(dd * dd, -0.72134752044448170368f, lg2g);
    float w = fast_ex2(fmaf(k.sigN, fast_lg2(dn), e));
    if (maskValue != q.w) w *= k.invM;
    sumC = Float3{fmaf(color.x, w, sumC.x), fmaf(color.y, w, sumC.y), fmaf(color.z, w, sumC.z)}; sumW += w;
#else
    const float w = bilateral_weight(n0, n1, d0, h2f(gg.w), maskValue != q.w, k, g);
    sumC = sumC + color * w; sumW += w;
#endif
}

// sum / weight: three IEEE divisions like the reference (RTRT_DN_FMA 0) or one reciprocal and three products
VD Float3 div_sum(Float3 c, float w) {
#if RTRT_DN_FMA
    return c * (1.0f / w);
#else
    return c / w;
#endif
}

template <int PAR, int TP>
VD void taps7(const ushort4* bC, const ushort4* bG, Float3 n0, float d0, unsigned short maskValue, const BilateralK& k, Float3& sumC, float& sumW) {
    const float fe = floor_eff(k);
#pragma unroll
    for (int i = 0; i < 24; ++i) {
        const int j = PAR + 2 * i, si = (j / 7) * TP + j % 7;
        tap_fast(bC[si], bG[si], n0, d0, maskValue, k, fe, kG7[j], kLg2G7[j], sumC, sumW);
    }
}

// -------------------------------------------------------------------------------------------------------------------
// Pass A: TemporalFilter (+ tile noise of its output).  16x16 pixels per CTA, 18x18 staged tile.
// -------------------------------------------------------------------------------------------------------------------
template <bool FILTER>
__global__ void __launch_bounds__(256, RTRT_DN_MINB) k_dn_temporal(const DnParams P, const ushort4* __restrict__ in, const ushort4* __restrict__ accum, ushort4* __restrict__ out) {
    constexpr int T = 18;
    using Stage = TileStage<T, 1>;
    constexpr int TP = Stage::TP, OX = Stage::OX;
    __shared__ __align__(128) ushort4 sC[T * TP];
    __shared__ __align__(128) ushort4 sG[T * TP];            // normal.xyz (half) + depth (half)
    __shared__ __align__(8) unsigned long long mbar;
    __shared__ ushort4 sOut[256];
    __shared__ unsigned short sDepth[256];
    __shared__ float sN8[4];
    const int tid = threadIdx.y * 16 + threadIdx.x;
    Stage::load(P, in, sC, sG, &mbar, tid, blockIdx.x, blockIdx.y + P.by0, 0, T, true);
    const int x = blockIdx.x * 16 + threadIdx.x, y = (blockIdx.y + P.by0) * 16 + threadIdx.y;
    const int ci = (threadIdx.y + 1) * TP + threadIdx.x + 1 + OX;
    ushort4 result = sC[ci];
    const ushort4 cg = sG[ci];
    if (FILTER && x < P.W && y < P.H) {
        const Float3 colorValue = px_rgb(result), normalValue = px_rgb(cg);
        const float depthValue = h2f(cg.w);
        const unsigned short maskValue = result.w;
        if (!(depthValue >= kRayMax)) {
            const BilateralK bk = bilateral_k(P.temp_n, P.temp_d, P.temp_m, 0.0f);
            Float3 nMax = rgb_to_ycocg(colorValue), nMin = nMax, filtered = f3(0.0f);
            float wsum = 0.0f;
            const float fe = floor_eff(bk);
#pragma unroll
            for (int j = 0; j < 9; ++j) {
                const int xo = j % 3, yo = j / 3;
                const int si = (threadIdx.y + yo) * TP + threadIdx.x + xo + OX;
                const ushort4 q = sC[si], g = sG[si];
                const Float3 color = px_rgb(q);
                tap_fast(q, g, normalValue, depthValue, maskValue, bk, fe, kG3[j], kLg2G3[j], filtered, wsum);
                const Float3 nc = rgb_to_ycocg(color);
                nMax = max3f(nMax, nc); nMin = min3f(nMin, nc);
            }
            if (wsum > 0.0f) filtered = div_sum(filtered, wsum); else filtered = f3(0.0f);
            if (isnan3(filtered)) filtered = f3(0.0f);
            const float2 mv = __half22float2(__ldg(P.motion + (size_t)y * P.W + x));
            const Float2 huv = (Float2{(float)x, (float)y} + 0.5f) * Float2{1.0f / (float)P.W, 1.0f / (float)P.H} + (Float2{mv.x, mv.y} - 0.5f);
            if (huv.x < 0 || huv.y < 0 || huv.x > 1.0f || huv.y > 1.0f) {
                result = px_pack(filtered, maskValue);
            } else {
                Float3 hist = sample_history(accum, P.aW, P.aH, huv);
                Float3 hy = rgb_to_ycocg(hist);
                hy = Float3{clampf(hy.x, nMin.x, nMax.x), clampf(hy.y, nMin.y, nMax.y), clampf(hy.z, nMin.z, nMax.z)};
                hist = ycocg_to_rgb(hy);
                const float lumaCurrent = rgb_to_ycocg(colorValue).x;
                float discard = 0.0f;
                const int hx = (int)floorf(huv.x * P.aW), hyy = (int)floorf(huv.y * P.aH);
#pragma unroll
                for (int i = 0; i < 4; ++i) discard += (maskValue != ld_px(accum, P.aW, P.aH, hx + (i & 1), hyy + (i >> 1)).w) ? 1.0f : 0.0f;
                discard /= 4.0f;
                hist = hist * (1.0f - discard) + filtered * discard;
                const float lumaHistory = rgb_to_ycocg(hist).x;
                if (isnan3(hist)) hist = f3(0.0f);
                Float3 o = temporal_blend(colorValue, hist, lumaHistory, nMin.x, nMax.x, lumaCurrent, 1.0f / 8.0f);
                if (isnan3(o)) o = f3(0.0f);
                result = px_pack(o, maskValue);
            }
        }
    }
    if (FILTER && x < P.W && y < P.H) out[(size_t)y * P.W + x] = result;
    sOut[tid] = result; sDepth[tid] = cg.w;
    __syncthreads();
    block_noise_levels(P, sOut, sDepth, sN8, tid);
}

// -------------------------------------------------------------------------------------------------------------------
// Pass B: SpatialFilter7x7 (24 of 49 taps, parity-alternating) + accumulation copy + tile noise of its output.
// 16x16 pixels per CTA, 22x22 staged tile.  gate16 = noise level of the INPUT (written by the previous pass).
// -------------------------------------------------------------------------------------------------------------------
template <bool FILTER>
__global__ void __launch_bounds__(256, RTRT_DN_MINB) k_dn_spatial7(const DnParams P, const ushort4* __restrict__ in, ushort4* __restrict__ out, ushort4* __restrict__ accumOut,
                                                     const __half* __restrict__ gate16, int wantNoise) {
    constexpr int T = 22;
    using Stage = TileStage<T, 3>;
    constexpr int TP = Stage::TP, OX = Stage::OX;
    __shared__ __align__(128) ushort4 sC[T * TP];
    __shared__ __align__(128) ushort4 sG[T * TP];
    __shared__ __align__(8) unsigned long long mbar;
    __shared__ ushort4 sOut[256];
    __shared__ unsigned short sDepth[256];
    __shared__ float sN8[4];
    const int tid = threadIdx.y * 16 + threadIdx.x;
    bool active = FILTER;
    if (FILTER) active = !(__half2float(__ldg(gate16 + (size_t)(blockIdx.y + P.by0) * ((P.W + 15) / 16) + blockIdx.x)) < P.thr_local);
    const int lo = active ? 0 : 3, hi = active ? T : T - 3;          // gated tiles only need their own 16x16 pixels
    Stage::load(P, in, sC, sG, &mbar, tid, blockIdx.x, blockIdx.y + P.by0, lo, hi, active);
    const int x = blockIdx.x * 16 + threadIdx.x, y = (blockIdx.y + P.by0) * 16 + threadIdx.y;
    const int ci = (threadIdx.y + 3) * TP + threadIdx.x + 3 + OX;
    ushort4 result = sC[ci];
    const ushort4 cg = sG[ci];
    if (active && x < P.W && y < P.H) {
        const Float3 normalValue = px_rgb(cg);
        const float depthValue = h2f(cg.w);
        const unsigned short maskValue = result.w;
        if (!(depthValue >= kRayMax)) {
            const BilateralK bk = bilateral_k(P.local_n, P.local_d, P.local_m, 0.0001f);
            Float3 sumC = f3(0.0f); float sumW = 0.0f;
#if RTRT_DN_FAST
            // the 24 taps of either parity with compile-time offsets: shared-memory addresses = one base + immediate, Gaussian = immediate
            const ushort4* bC = sC + threadIdx.y * TP + threadIdx.x + OX;
            const ushort4* bG = sG + threadIdx.y * TP + threadIdx.x + OX;
            if (P.frameNum % 2) taps7<1, TP>(bC, bG, normalValue, depthValue, maskValue, bk, sumC, sumW);
            else taps7<0, TP>(bC, bG, normalValue, depthValue, maskValue, bk, sumC, sumW);
#else
            int j = P.frameNum % 2;
#pragma unroll 4
            for (int i = 0; i < 24; ++i) {
                const int xo = j % 7, yo = j / 7;
                const int si = (threadIdx.y + yo) * TP + threadIdx.x + xo + OX;
                const ushort4 q = sC[si], g = sG[si];
                const float w = bilateral_weight(normalValue, px_rgb(g), depthValue, h2f(g.w), maskValue != q.w, bk, cG7[j]);
                sumC = sumC + px_rgb(q) * w; sumW += w;
                j += 2;
            }
#endif
            if (isnan3(sumC)) sumC = f3(0.0f);
            if (sumW != sumW) sumW = 0.0f;
            Float3 fin = (sumW == 0.0f) ? f3(0.0f) : div_sum(sumC, sumW);
            if (isnan3(fin)) fin = f3(0.0f);
            result = px_pack(fin, maskValue);
        }
    }
    if (x < P.W && y < P.H) {
        if (FILTER) out[(size_t)y * P.W + x] = result;
        if (accumOut) accumOut[(size_t)y * P.W + x] = result;        // CopyToHistoryColorBuffer (:159-170)
    }
    if (wantNoise) {
        sOut[tid] = result; sDepth[tid] = cg.w;
        __syncthreads();
        block_noise_levels(P, sOut, sDepth, sN8, tid);
    }
}

// -------------------------------------------------------------------------------------------------------------------
// Pass C: SpatialFilterGlobal5x5<S>, S = 3, 6, 12 — dilated taps straight from global memory (L1/L2 resident).
// -------------------------------------------------------------------------------------------------------------------
// one tap exactly as the reference loop orders it (src/temporalDenoising.cuh:556-582): NaN colours drop out with their weight
VD void tap5_exact(ushort4 q, ushort4 gg, Float3 normalValue, float depthValue, unsigned short maskValue, const BilateralK& bk, float g, Float3& sumC, float& sumW) {
    Float3 color = px_rgb(q);
    float w = bilateral_weight(normalValue, px_rgb(gg), depthValue, h2f(gg.w), maskValue != q.w, bk, g);
    if (isnan3(color)) { color = f3(0.0f); w = 0.0f; }
    sumC = sumC + color * w; sumW += w;
}

template <int S>
__global__ void __launch_bounds__(256, RTRT_DN_MINB) k_dn_spatial5(const DnParams P, const ushort4* __restrict__ in, ushort4* __restrict__ out, const __half* __restrict__ gate16) {
    const int x = blockIdx.x * 16 + threadIdx.x, y = (blockIdx.y + P.by0) * 16 + threadIdx.y;
    if (x >= P.W || y >= P.H) return;
    const size_t c = (size_t)y * P.W + x;
    ushort4 result = __ldg(in + c);
    const bool active = !(__half2float(__ldg(gate16 + (size_t)(blockIdx.y + P.by0) * ((P.W + 15) / 16) + blockIdx.x)) < P.thr_large);
    if (active) {
        const ushort4 cg = __ldg(P.guide + c);
        const Float3 normalValue = px_rgb(cg);
        const float depthValue = h2f(cg.w);
        const unsigned short maskValue = result.w;
        if (!(depthValue >= 10e9f)) {
            const BilateralK bk = bilateral_k(P.large_n, P.large_d, P.large_m, 0.0f);
            Float3 sumC = f3(0.0f); float sumW = 0.0f;
            // uniform per CTA: every tap of every pixel of the tile is inside the image
            const int tx0 = blockIdx.x * 16, ty0 = (blockIdx.y + P.by0) * 16;
            const bool interior = RTRT_DN_FAST && tx0 >= 2 * S && tx0 + 15 + 2 * S < P.W && ty0 >= 2 * S && ty0 + 15 + 2 * S < P.H;
            bool redo = !interior;
            if (interior) {
                // rows outside, columns inside, both fully unrolled: the column offsets and the Gaussian weights are immediates, one pointer
                // step per row and image (same tap order k = 5 j + i as the reference loop)
                const ptrdiff_t rowStep = (ptrdiff_t)S * P.W;
                const ushort4* rc = in + c - 2 * rowStep;
                const ushort4* rg = P.guide + c - 2 * rowStep;
                const float fe = floor_eff(bk);
#pragma unroll
                for (int j = 0; j < 5; ++j) {
#pragma unroll
                    for (int i = 0; i < 5; ++i)
                        tap_fast(__ldg(rc + (i - 2) * S), __ldg(rg + (i - 2) * S), normalValue, depthValue, maskValue, bk, fe, kG5[5 * j + i], kLg2G5[5 * j + i], sumC, sumW);
                    rc += rowStep; rg += rowStep;
                }
                redo = isnan3(sumC) || sumW != sumW;          // a NaN colour (or weight) somewhere: the exact loop decides what drops out
            }
            if (redo) {
                sumC = f3(0.0f); sumW = 0.0f;
#pragma unroll 1
                for (int j = 0; j < 5; ++j) {
                    const int sy = clampi(y + (j - 2) * S, 0, P.H - 1);
                    const size_t row = (size_t)sy * P.W;
#pragma unroll 1
                    for (int i = 0; i < 5; ++i) {
                        const int sx = clampi(x + (i - 2) * S, 0, P.W - 1);
                        const size_t g = row + sx;
                        tap5_exact(__ldg(in + g), __ldg(P.guide + g), normalValue, depthValue, maskValue, bk, cG5[5 * j + i], sumC, sumW);
                    }
                }
            }
            if (isnan3(sumC)) sumC = f3(0.0f);
            if (sumW != sumW) sumW = 0.0f;
            Float3 fin = (sumW == 0.0f) ? f3(0.0f) : div_sum(sumC, sumW);
            if (isnan3(fin)) fin = f3(0.0f);
            result = px_pack(fin, maskValue);
        }
    }
    out[c] = result;
}

// -------------------------------------------------------------------------------------------------------------------
// Pass D: ApplyAlbedo (:1127-1139) + TemporalFilter2 (:896-1111) + CopyToHistoryColorDepthBuffer (:142-157).
// The remodulated colour (half-rounded, mask channel = half(1.0) like the reference's Store2DHalf4) is staged in an
// 18x18 tile as YCoCg; the history of the previous frame is read from histPrev, this frame's goes to histNext.
// -------------------------------------------------------------------------------------------------------------------
template <bool TF2>
__global__ void __launch_bounds__(256, RTRT_DN_MINB) k_dn_final(const DnParams P, const ushort4* __restrict__ in, ushort4* __restrict__ out,
                                                  const ushort4* __restrict__ histPrev, ushort4* __restrict__ histNext, __half* __restrict__ histDepth) {
    constexpr int T = 18;
    __shared__ Float3 sY[T * T];
    const int tid = threadIdx.y * 16 + threadIdx.x;
    const int x = blockIdx.x * 16 + threadIdx.x, y = (blockIdx.y + P.by0) * 16 + threadIdx.y;
    const unsigned short kOne = 0x3C00;                                   // half(1.0): what ApplyAlbedo leaves in the mask channel
    auto remodulated = [&](int gx, int gy) {
        const size_t g = (size_t)clampi(gy, 0, P.H - 1) * P.W + clampi(gx, 0, P.W - 1);
        const Float3 c = px_rgb(__ldg(in + g)) * px_rgb(__ldg(P.albedo + g));
        return px_pack(c, kOne);
    };
    ushort4 result;
    if (TF2) {
        const int x0 = blockIdx.x * 16 - 1, y0 = (blockIdx.y + P.by0) * 16 - 1;
        for (int i = tid; i < T * T; i += 256) sY[i] = rgb_to_ycocg(px_rgb(remodulated(x0 + i % T, y0 + i / T)));
        __syncthreads();
        if (x >= P.W || y >= P.H) return;
        const Float3 centerY = sY[(threadIdx.y + 1) * T + threadIdx.x + 1];
        const Float3 colorValue = ycocg_to_rgb(centerY);
        result = remodulated(x, y);
        // all masks equal half(1.0) after ApplyAlbedo, so every tap takes part in the min/max
        Float3 nMax = f3(1.175494351e-38f), nMin = f3(3.402823466e+38f), nMax2 = nMax, nMin2 = nMin;     // Float3(FLT_MIN) / Float3(FLT_MAX), :968-972
#pragma unroll
        for (int j = 0; j < 9; ++j) {
            const int xo = j % 3, yo = j / 3;
            const Float3 color = sY[(threadIdx.y + yo) * T + threadIdx.x + xo];
            nMax = max3f(nMax, color); nMin = min3f(nMin, color);
            if (abs(xo - 1) + abs(yo - 1) <= 1) { nMax2 = max3f(nMax2, color); nMin2 = min3f(nMin2, color); }
        }
        nMax = (nMax + nMax2) / 2.0f; nMin = (nMin + nMin2) / 2.0f;
        const float2 mv = __half22float2(__ldg(P.motion + (size_t)y * P.W + x));
        const Float2 huv = (Float2{(float)x, (float)y} + 0.5f) * Float2{1.0f / (float)P.W, 1.0f / (float)P.H} + (Float2{mv.x, mv.y} - 0.5f);
        if (!(huv.x < 0 || huv.y < 0 || huv.x > 1.0f || huv.y > 1.0f)) {
            Float3 ch = sample_history(histPrev, P.hW, P.hH, huv);
            Float3 hy = rgb_to_ycocg(ch);
            hy = Float3{clampf(hy.x, nMin.x, nMax.x), clampf(hy.y, nMin.y, nMax.y), clampf(hy.z, nMin.z, nMax.z)};
            ch = ycocg_to_rgb(hy);
            const float lumaCurrent = rgb_to_ycocg(colorValue).x;
            float discard = 0.0f;
            const int hx = (int)floorf(huv.x * P.hW), hyy = (int)floorf(huv.y * P.hH);
#pragma unroll
            for (int i = 0; i < 4; ++i) discard += (kOne != ld_px(histPrev, P.hW, P.hH, hx + (i & 1), hyy + (i >> 1)).w) ? 1.0f : 0.0f;
            discard /= 4.0f;
            if (discard != 1.0f) {
                ch = ch * (1.0f - discard) + colorValue * discard;
                const float lumaHistory = rgb_to_ycocg(ch).x;
                Float3 o = temporal_blend(colorValue, ch, lumaHistory, nMin.x, nMax.x, lumaCurrent, 3.0f / 4.0f);
                if (isnan3(o)) o = f3(0.0f);
                result = px_pack(o, kOne);
            }
        }
    } else {
        if (x >= P.W || y >= P.H) return;
        result = remodulated(x, y);
    }
    const size_t c = (size_t)y * P.W + x;
    out[c] = result;
    if (histNext) { histNext[c] = result; histDepth[c] = __ldg(P.depth + c); }
}

// -------------------------------------------------------------------------------------------------------------------
// Progressive accumulation (BASELINE config 5: N spp without the denoiser): sum += colour * albedo of the frame just path
// traced (the remodulation of ApplyAlbedo, src/temporalDenoising.cuh:1127-1139, in fp32), sample count in .w; rows
// [rowBegin, rowEnd) only, so row-sharded GPUs accumulate their own band.
// -------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_accumulate(const ushort4* __restrict__ color, const ushort4* __restrict__ albedo, float4* __restrict__ acc,
                                                    int W, int localRows, int rowBegin, int stripH, int stripShards, int stripShard, int reset) {
    const size_t k = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (k >= (size_t)W * localRows) return;
    const int j = (int)(k / W), x = (int)(k - (size_t)j * W);
    int y = rowBegin + j;
    if (stripShards > 1) { const int q = j / stripH; y = (q * stripShards + stripShard) * stripH + (j - q * stripH); }
    const size_t i = (size_t)y * W + x;
    const Float3 c = px_rgb(__ldg(color + i)) * px_rgb(__ldg(albedo + i));
    float4 a = reset ? make_float4(0.0f, 0.0f, 0.0f, 0.0f) : acc[i];
    a.x += c.x; a.y += c.y; a.z += c.z; a.w += 1.0f;
    acc[i] = a;
}

int launch_accumulate(Context& c, int reset) {
    if (!c.dColor || c.gbufW != c.width || c.gbufH != c.height) return c.fail(RTRT_ERR_STATE, "rtrt_accumulate: no path-traced frame");
    const size_t px = (size_t)c.width * c.height;
    if (!c.dAccumRadiance || c.accumRadiancePx != px) {
        if (c.dAccumRadiance) cudaFree(c.dAccumRadiance);
        c.dAccumRadiance = nullptr; c.accumRadiancePx = 0;
        RTRT_CUDA(c, cudaMalloc(&c.dAccumRadiance, px * sizeof(float4)));
        RTRT_CUDA(c, cudaMemsetAsync(c.dAccumRadiance, 0, px * sizeof(float4), c.stream));
        c.accumRadiancePx = px;
    }
    const int rows = local_rows(c);
    const size_t n = (size_t)rows * c.width;
    k_accumulate<<<(unsigned)((n + 255) / 256), 256, 0, c.stream>>>(reinterpret_cast<const ushort4*>(c.dColor), reinterpret_cast<const ushort4*>(c.dAlbedo),
                                                                   reinterpret_cast<float4*>(c.dAccumRadiance), c.width, rows, c.rowBegin, c.stripH,
                                                                   c.stripShards, c.stripShard, reset);
    c.launches++;
    RTRT_CUDA(c, cudaGetLastError());
    return RTRT_OK;
}

// -------------------------------------------------------------------------------------------------------------------
int launch_denoise(Context& c) {
    if (!c.dColor || c.gbufW != c.width || c.gbufH != c.height) return c.fail(RTRT_ERR_STATE, "rtrt_denoise: no path-traced frame to denoise");
    DnParams P{};
    P.W = c.width; P.H = c.height; P.frameNum = c.frameNum;
    P.aW = c.accumW ? c.accumW : c.width; P.aH = c.accumH ? c.accumH : c.height;
    P.hW = c.histW ? c.histW : c.width; P.hH = c.histH ? c.histH : c.height;
    const rtrt_denoise_params& d = c.denoiseParams;
    P.local_n = d.local_denoise_sigma_normal; P.local_d = d.local_denoise_sigma_depth; P.local_m = d.local_denoise_sigma_material;
    P.large_n = d.large_denoise_sigma_normal; P.large_d = d.large_denoise_sigma_depth; P.large_m = d.large_denoise_sigma_material;
    P.temp_n = d.temporal_denoise_sigma_normal; P.temp_d = d.temporal_denoise_sigma_depth; P.temp_m = d.temporal_denoise_sigma_material;
    P.thr_local = d.noise_threshold_local; P.thr_large = d.noise_threshold_large;
    P.normal = reinterpret_cast<const ushort4*>(c.dNormal); P.depth = reinterpret_cast<const __half*>(c.dDepth);
    P.motion = reinterpret_cast<const __half2*>(c.dMotion); P.albedo = reinterpret_cast<const ushort4*>(c.dAlbedo);
    P.noise8 = reinterpret_cast<__half*>(c.dNoise); P.noise16 = reinterpret_cast<__half*>(c.dNoise16);
    P.guide = reinterpret_cast<const ushort4*>(c.dGuide);
    const rtrt_pass_settings& ps = c.passSettings;
    const bool temporal = ps.enableTemporalDenoising != 0, local = ps.enableLocalSpatialFilter != 0, wide = ps.enableWideSpatialFilter != 0,
               tf2 = ps.enableTemporalDenoising2 != 0, notFirst = c.frameNum != 1;
    // Row-band mode (multi-GPU, rtrt_set_row_range): only the 16-row block rows that can influence rows [rowBegin, rowEnd) are
    // filtered.  Dependency reach of the band through the passes, walking backwards: TemporalFilter2 1 row, the dilated 5x5
    // passes 24 + 12 + 6, their 16x16 noise gate (tile-aligned, so up to 15 more rows of SpatialFilter7x7 output), 7x7: 3,
    // its own gate 15, TemporalFilter 1  => 77 rows; kBandHalo = 80 = five block rows.  Every pass covers the same block rows:
    // what the outer halo rows compute from rows outside is garbage by construction and never reaches the band.
    // The caller must have made ALL rows of the G-buffers and history valid first (all-gather of the bands).
    // With the 5-iteration schedule (enableWideSpatialFilter == 2: strides 1, 2, 4, 8, 16) the dilated passes reach 62 rows instead of 42
    // => 97 rows; the halo is then seven block rows.
    const bool sched5 = ps.enableWideSpatialFilter == 2;
    const bool band = c.rowBegin > 0 || c.rowEnd < c.height;
    const int kBandHalo = c.bandHaloRows >= 0 ? c.bandHaloRows : (sched5 ? 112 : 80);     // 0 with the multi-GPU halo exchange (multi.cu)
    // halo exchange: before a pass that reads `rows` rows beyond the band, the multi-GPU driver fetches them from the neighbouring bands
    auto halo = [&](ushort4* img, int rows) -> int { return (c.haloFn && band) ? c.haloFn(c, img, rows) : RTRT_OK; };
    const int blocksY = (c.height + 15) / 16;
    const int by0 = band ? (c.rowBegin - kBandHalo < 0 ? 0 : (c.rowBegin - kBandHalo) / 16) : 0;
    const int by1 = band ? (c.rowEnd + kBandHalo + 15) / 16 < blocksY ? (c.rowEnd + kBandHalo + 15) / 16 : blocksY : blocksY;
    P.by0 = by0;
    const dim3 grid((c.width + 15) / 16, by1 - by0), block(16, 16);
    ushort4* cur = reinterpret_cast<ushort4*>(c.dColor); ushort4* oth = reinterpret_cast<ushort4*>(c.dColorTmp);
    ushort4* accum = reinterpret_cast<ushort4*>(c.dAccum);
    const __half* n16 = reinterpret_cast<const __half*>(c.dNoise16);
    auto swap = [&]() { ushort4* t = cur; cur = oth; oth = t; };
    const size_t bytes = (size_t)c.width * c.height * sizeof(ushort4);

    {
        // guide rows: the filtered block rows + the 24-row reach of the widest dilated pass
        const int gr = sched5 ? 32 : 24;
        const int g0 = by0 * 16 - gr < 0 ? 0 : by0 * 16 - gr, g1 = by1 * 16 + gr > c.height ? c.height : by1 * 16 + gr;
        const size_t first = (size_t)g0 * c.width, n = (size_t)(g1 - g0) * c.width;
        k_dn_pack_guide<<<(unsigned)((n + 255) / 256), 256, 0, c.stream>>>(P.normal, P.depth, reinterpret_cast<ushort4*>(c.dGuide), first, n);
        c.launches++;
    }
    // TemporalFilter (src/denoising.cu:49-64) [+ noise level of its output, :68-76]
    if (temporal && notFirst) { k_dn_temporal<true><<<grid, block, 0, c.stream>>>(P, cur, accum, oth); swap(); c.launches++; }
    else if (local) { k_dn_temporal<false><<<grid, block, 0, c.stream>>>(P, cur, accum, oth); c.launches++; }
    // SpatialFilter7x7 (:90-98) + CopyToHistoryColorBuffer (:100-106) [+ noise level for the wide passes, :110-118]
    if (local) {
        { int rc = halo(cur, 3); if (rc) return rc; }
        k_dn_spatial7<true><<<grid, block, 0, c.stream>>>(P, cur, oth, temporal ? accum : nullptr, n16, wide ? 1 : 0); swap(); c.launches++;
    }
    else if (temporal || wide) { k_dn_spatial7<false><<<grid, block, 0, c.stream>>>(P, cur, oth, temporal ? accum : nullptr, n16, wide ? 1 : 0); c.launches++; }
    // SpatialFilterGlobal5x5<3|6|12> (:132-157)
    if (wide && sched5) {                                     // BASELINE config 4: five a-trous iterations, strides 1, 2, 4, 8, 16
        { int rc = halo(cur, 2); if (rc) return rc; }
        k_dn_spatial5<1><<<grid, block, 0, c.stream>>>(P, cur, oth, n16); swap();
        { int rc = halo(cur, 4); if (rc) return rc; }
        k_dn_spatial5<2><<<grid, block, 0, c.stream>>>(P, cur, oth, n16); swap();
        { int rc = halo(cur, 8); if (rc) return rc; }
        k_dn_spatial5<4><<<grid, block, 0, c.stream>>>(P, cur, oth, n16); swap();
        { int rc = halo(cur, 16); if (rc) return rc; }
        k_dn_spatial5<8><<<grid, block, 0, c.stream>>>(P, cur, oth, n16); swap();
        { int rc = halo(cur, 32); if (rc) return rc; }
        k_dn_spatial5<16><<<grid, block, 0, c.stream>>>(P, cur, oth, n16); swap();
        c.launches += 5;
    } else if (wide) {
        { int rc = halo(cur, 6); if (rc) return rc; }
        k_dn_spatial5<3><<<grid, block, 0, c.stream>>>(P, cur, oth, n16); swap();
        { int rc = halo(cur, 12); if (rc) return rc; }
        k_dn_spatial5<6><<<grid, block, 0, c.stream>>>(P, cur, oth, n16); swap();
        { int rc = halo(cur, 24); if (rc) return rc; }
        k_dn_spatial5<12><<<grid, block, 0, c.stream>>>(P, cur, oth, n16); swap();
        c.launches += 3;
    }
    // ApplyAlbedo (:160-163) + TemporalFilter2 (:165-179) + CopyToHistoryColorDepthBuffer (:180-186)
    ushort4* histPrev = reinterpret_cast<ushort4*>(c.dHistoryColor); ushort4* histNext = reinterpret_cast<ushort4*>(c.dHistoryAlbedo);
    if (tf2 && notFirst) {
        { int rc = halo(cur, 1); if (rc) return rc; }
        k_dn_final<true><<<grid, block, 0, c.stream>>>(P, cur, oth, histPrev, histNext, reinterpret_cast<__half*>(c.dHistoryDepth));
    }
    else k_dn_final<false><<<grid, block, 0, c.stream>>>(P, cur, oth, histPrev, tf2 ? histNext : nullptr, reinterpret_cast<__half*>(c.dHistoryDepth));
    swap(); c.launches++;
    // band mode with all-gathered G-buffers: the history depth of the rows outside the band (k_dn_final stores the band's).  With the halo exchange the
    // depth is only valid around the band and RTRT_BUF_HISTORY_DEPTH (an observation point: no kernel reads it) stays band-local.
    if (band && tf2 && !c.haloFn) RTRT_CUDA(c, cudaMemcpyAsync(c.dHistoryDepth, c.dDepth, (size_t)c.width * c.height * sizeof(__half), cudaMemcpyDeviceToDevice, c.stream));
    if (temporal) { c.accumW = c.width; c.accumH = c.height; }
    if (tf2) { c.histW = c.width; c.histH = c.height; }
    if (tf2) { void* t = c.dHistoryColor; c.dHistoryColor = c.dHistoryAlbedo; c.dHistoryAlbedo = t; }       // history ping-pong
    (void)bytes;
    c.dColor = cur; c.dColorTmp = oth;          // RTRT_BUF_RENDER_COLOR now names the denoised frame
    RTRT_CUDA(c, cudaGetLastError());
    return RTRT_OK;
}

}  // namespace rtrt
