// rtrt_microbench — the two ceilings a BVH traversal without RT cores actually runs against on this GPU, measured with the library's
// own kernels so that bench.py can state a denominator the traversal kernel could reach ("the bench's roofline is not a
// roofline" — node fetches are L1/L2 hits, HBM bandwidth is the wrong peak):
//   * L2 read bandwidth: every SM streams a 32 MiB buffer (resident in the 126 MB L2 after the first pass) with 16-byte ld.global.cg
//     loads (L1 bypassed), all SMs, enough CTAs to saturate the crossbar;
//   * dependent-load latency: ONE thread chases a random cyclic permutation (one 128-byte line per element) through
//       64 KiB   -> L1 hit latency      (ld.global.ca, what a node re-visit costs)
//       16 MiB   -> L2 hit latency      (ld.global.cg)
//       512 MiB  -> DRAM latency        (ld.global.cg, working set 4x the L2)
//     which is the cost of ONE step of the pointer chase "pop node -> fetch -> test -> next child" before any arithmetic.
// Runs outside every timed region; allocates and frees its own buffers.
#include "context.h"
#include <vector>
#include <random>
#include <algorithm>

namespace rtrt {

__device__ __forceinline__ uint4 ld_cg_volatile(const uint4* p) {      // volatile: the pass loop must re-issue every load
    uint4 v;
    asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

__global__ void __launch_bounds__(256) k_mb_l2_read(const uint4* __restrict__ buf, size_t n16, int passes, unsigned* sink) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    unsigned acc = 0u;
    for (int p = 0; p < passes; ++p) {
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += 4 * stride) {
            // four independent 16-byte loads in flight per thread
            uint4 a = ld_cg_volatile(buf + i), b = make_uint4(0, 0, 0, 0), c = b, d = b;
            if (i + stride < n16) b = ld_cg_volatile(buf + i + stride);
            if (i + 2 * stride < n16) c = ld_cg_volatile(buf + i + 2 * stride);
            if (i + 3 * stride < n16) d = ld_cg_volatile(buf + i + 3 * stride);
            acc += a.x ^ b.y ^ c.z ^ d.w;
        }
    }
    if (acc == 0x12345678u) *sink = acc;       // keeps the loads alive
}

template <bool L1>
__global__ void k_mb_chase(const uint32_t* __restrict__ next, uint32_t start, int steps, uint32_t* out, long long* cycles) {
    uint32_t i = start;
    // warm: one full lap is done by the host-chosen `steps` of an untimed launch
    const long long t0 = clock64();
    for (int s = 0; s < steps; ++s) {
        i = L1 ? __ldca(next + (size_t)i * 32) : __ldcg(next + (size_t)i * 32);
    }
    const long long t1 = clock64();
    *out = i; *cycles = t1 - t0;
}

static int chase(Context& c, size_t bytes, bool l1, int steps, float* nsOut, float* cycOut) {
    const size_t lines = bytes / 128;
    std::vector<uint32_t> perm(lines), next(lines * 32, 0u);
    for (size_t i = 0; i < lines; ++i) perm[i] = (uint32_t)i;
    std::mt19937 rng(12345);
    std::shuffle(perm.begin(), perm.end(), rng);
    for (size_t i = 0; i < lines; ++i) next[(size_t)perm[i] * 32] = perm[(i + 1) % lines];      // one cycle through every line
    uint32_t* dNext = nullptr; uint32_t* dOut = nullptr; long long* dCyc = nullptr;
    RTRT_CUDA(c, cudaMalloc((void**)&dNext, lines * 128));
    RTRT_CUDA(c, cudaMalloc((void**)&dOut, 4)); RTRT_CUDA(c, cudaMalloc((void**)&dCyc, 8));
    RTRT_CUDA(c, cudaMemcpy(dNext, next.data(), lines * 128, cudaMemcpyHostToDevice));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int warm = (int)std::min<size_t>(lines, 1u << 17);            // pull the working set into the cache level it fits
    float best = 1e30f; long long bestCyc = 0;
    for (int rep = 0; rep < 3; ++rep) {
        if (l1) k_mb_chase<true><<<1, 1, 0, c.stream>>>(dNext, perm[0], warm, dOut, dCyc); else k_mb_chase<false><<<1, 1, 0, c.stream>>>(dNext, perm[0], warm, dOut, dCyc);
        cudaEventRecord(e0, c.stream);
        if (l1) k_mb_chase<true><<<1, 1, 0, c.stream>>>(dNext, perm[0], steps, dOut, dCyc); else k_mb_chase<false><<<1, 1, 0, c.stream>>>(dNext, perm[0], steps, dOut, dCyc);
        cudaEventRecord(e1, c.stream);
        RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
        float ms = 0.0f; cudaEventElapsedTime(&ms, e0, e1);
        long long cyc = 0; cudaMemcpy(&cyc, dCyc, 8, cudaMemcpyDeviceToHost);
        if (ms < best) { best = ms; bestCyc = cyc; }
    }
    c.launches += 6;
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(dNext); cudaFree(dOut); cudaFree(dCyc);
    *nsOut = best * 1e6f / (float)steps;
    *cycOut = (float)bestCyc / (float)steps;
    return RTRT_OK;
}

}  // namespace rtrt

using namespace rtrt;

extern "C" int rtrt_microbench(rtrt_ctx* ctx, rtrt_microbench_result* out) {
    if (!ctx) return RTRT_ERR_ARG;
    Context& c = ctx->c;
    if (!out) return c.fail(RTRT_ERR_ARG, "rtrt_microbench: null out");
    *out = rtrt_microbench_result{};
    RTRT_CUDA(c, cudaSetDevice(c.device));
    RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    {   // L2 read bandwidth
        const size_t bytes = 32u << 20, n16 = bytes / 16;
        uint4* buf = nullptr; unsigned* sink = nullptr;
        RTRT_CUDA(c, cudaMalloc((void**)&buf, bytes)); RTRT_CUDA(c, cudaMalloc((void**)&sink, 4));
        RTRT_CUDA(c, cudaMemsetAsync(buf, 1, bytes, c.stream));
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        const int grid = c.smCount * 8, passes = 20;
        float best = 1e30f;
        for (int rep = 0; rep < 4; ++rep) {
            k_mb_l2_read<<<grid, 256, 0, c.stream>>>(buf, n16, 1, sink);              // warm: buffer into L2
            cudaEventRecord(e0, c.stream);
            k_mb_l2_read<<<grid, 256, 0, c.stream>>>(buf, n16, passes, sink);
            cudaEventRecord(e1, c.stream);
            RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
            float ms = 0.0f; cudaEventElapsedTime(&ms, e0, e1);
            best = ms < best ? ms : best;
        }
        c.launches += 8;
        out->l2_read_gbs = (float)((double)bytes * passes / (best * 1e-3) / 1e9);
        cudaEventDestroy(e0); cudaEventDestroy(e1);
        cudaFree(buf); cudaFree(sink);
    }
    int rc;
    if ((rc = chase(c, 64u << 10, true, 1 << 16, &out->l1_latency_ns, &out->l1_latency_cycles))) return rc;
    if ((rc = chase(c, 16u << 20, false, 1 << 16, &out->l2_latency_ns, &out->l2_latency_cycles))) return rc;
    if ((rc = chase(c, (size_t)512 << 20, false, 1 << 15, &out->dram_latency_ns, &out->dram_latency_cycles))) return rc;
    return RTRT_OK;
}
