// rtrt_traverse_host_noalloc — the allocation-free variant of rtrt_traverse_host (capi.cu): same closest-hit / any-hit query
// through HOST buffers, but the device staging of rays and hits and the visit counters live in ONE file-scope __device__ array
// instead of cudaMalloc'ed scratch, for callers that must not allocate (CUDA graph capture around their own work, tight serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
//
// Rules of the static scratch:
//   * every GPU that loads librtrt.so holds kNoAllocBytes (3 MiB + 256 B) of it until the process exits;
//   * in HOST code `gNoAllocScratch` is only a stand-in for the symbol, not a device address: the address is taken with
//     cudaGetSymbolAddress, once per device, with that device current, and that pointer is what the copies and the kernel get;
//   * all calls (all contexts on a device) share the array and it keeps the last call's contents: the counters are cleared by
//     every call, the ray / hit regions are fully rewritten for [0, n) before they are read; calls that overlap on one GPU race —
//     this entry point synchronises its stream before returning, so calls from one host thread never overlap;
//   * nothing here calls cudaMalloc / cudaFree or a routine that does: the traversal always uses the reference-layout arrays of
//     the last rtrt_build_bvh (the wide layout may have to be (re)built, which allocates — option wide_bvh is ignored here).
#include "traverse.cuh"
#include "context.h"
#include <mutex>

namespace rtrt {

constexpr uint32_t kNoAllocMaxRays = RTRT_NOALLOC_MAX_RAYS;
constexpr size_t kNoAllocCounterBytes = 256;
constexpr size_t kNoAllocRayOff = kNoAllocCounterBytes;
constexpr size_t kNoAllocHitOff = kNoAllocRayOff + (size_t)kNoAllocMaxRays * sizeof(Ray);
constexpr size_t kNoAllocBytes = kNoAllocHitOff + (size_t)kNoAllocMaxRays * sizeof(Hit);
static_assert(sizeof(Ray) == 32 && sizeof(Hit) == 16, "staging layout");

static __device__ __align__(256) unsigned char gNoAllocScratch[kNoAllocBytes];

constexpr int kNaThreads = 128;
constexpr int kNaMaxDevices = 64;
static std::mutex sNaMutex;
static unsigned char* sNaPtr[kNaMaxDevices] = {};          // device address of gNoAllocScratch on device d (0 = not looked up yet)

// k_traverse (traverse.cu) with its inputs, outputs and counters in the static scratch: scratch = the device address of
// gNoAllocScratch; counters[0] += node visits, counters[1] += leaf tests (cleared by the host before every launch).
template <int STACK>
__global__ void __launch_bounds__(kNaThreads) k_traverse_static(SceneView sc, unsigned char* scratch, uint32_t n, int anyHit) {
    __shared__ uint32_t sLink[STACK * kNaThreads];
    __shared__ float sT[STACK * kNaThreads];
    const uint32_t i = blockIdx.x * kNaThreads + threadIdx.x;
    uint32_t nodeVisits = 0, leafTests = 0;
    if (i < n) {
        const float4 a = __ldg(sc.tlasRows), b = __ldg(sc.tlasRows + 1), c = __ldg(sc.tlasRows + 2);     // scene box, src/traverse.h:127-129
        AABB scene;
        scene.max = Float3{sel_max(a.x, c.x), sel_max(a.y, c.y), sel_max(a.z, c.z)};
        scene.min = Float3{sel_min(b.x, a.w), sel_min(b.y, b.w), sel_min(b.z, c.w)};
        const float4* r4 = reinterpret_cast<const float4*>(scratch + kNoAllocRayOff) + 2 * (size_t)i;
        const float4 o = r4[0], d = r4[1];                   // plain loads: the scratch is rewritten between launches
        const RayPrep rp = prepare_ray(Float3{o.x, o.y, o.z}, Float3{d.x, d.y, d.z}, scene);
        const TraverseResult res = traverse_bvh<STACK, true>(sc, rp, anyHit, sLink + threadIdx.x, sT + threadIdx.x, kNaThreads);
        reinterpret_cast<float4*>(scratch + kNoAllocHitOff)[i] = make_float4(res.t, res.u, res.v, __int_as_float(res.tri));
        nodeVisits = res.nodeVisits; leafTests = res.leafTests;
    }
    for (int o = 16; o > 0; o >>= 1) { nodeVisits += __shfl_xor_sync(0xffffffffu, nodeVisits, o); leafTests += __shfl_xor_sync(0xffffffffu, leafTests, o); }
    if ((threadIdx.x & 31) == 0 && (nodeVisits | leafTests)) {
        unsigned long long* counters = reinterpret_cast<unsigned long long*>(scratch);
        atomicAdd(counters, (unsigned long long)nodeVisits); atomicAdd(counters + 1, (unsigned long long)leafTests);
    }
}

SceneView make_scene_view(const Context& c);   // traverse.cu

// set-up, once per device, with that device current
static int noalloc_scratch(Context& c, unsigned char** out) {
    if (c.device < 0 || c.device >= kNaMaxDevices) return c.fail(RTRT_ERR_CAPACITY, "rtrt_traverse_host_noalloc: device index above 63");
    std::lock_guard<std::mutex> lock(sNaMutex);
    if (!sNaPtr[c.device]) {
        unsigned char* p = nullptr;
        RTRT_CUDA(c, cudaGetSymbolAddress((void**)&p, gNoAllocScratch));
        if (!p) return c.fail(RTRT_ERR_CUDA, "rtrt_traverse_host_noalloc: cudaGetSymbolAddress returned a null address");
        sNaPtr[c.device] = p;
    }
    *out = sNaPtr[c.device];
    return RTRT_OK;
}

}  // namespace rtrt

using namespace rtrt;

extern "C" int rtrt_traverse_host_noalloc(rtrt_ctx* ctx, const rtrt_ray* h_rays, uint32_t n, rtrt_hit* h_hits, int any_hit,
                                          uint64_t* node_visits, uint64_t* leaf_tests) {
    if (!ctx) return RTRT_ERR_ARG;
    Context& c = ctx->c;
    if ((!h_rays || !h_hits) && n) return c.fail(RTRT_ERR_ARG, "rtrt_traverse_host_noalloc: null buffer");
    if (n > kNoAllocMaxRays) {
        char msg[160];
        snprintf(msg, sizeof(msg), "rtrt_traverse_host_noalloc: %u rays exceed the static scratch (RTRT_NOALLOC_MAX_RAYS = %u); split the batch or use rtrt_traverse_host",
                 n, kNoAllocMaxRays);
        return c.fail(RTRT_ERR_CAPACITY, msg);
    }
    if (c.stackSize > 32) return c.fail(RTRT_ERR_ARG, "rtrt_traverse_host_noalloc: stack_size must be <= 32");
    if (!c.bvhValid) return c.fail(RTRT_ERR_STATE, "rtrt_traverse_host_noalloc: BVH not built");
    if (node_visits) *node_visits = 0;
    if (leaf_tests) *leaf_tests = 0;
    if (n == 0) return RTRT_OK;
    RTRT_CUDA(c, cudaSetDevice(c.device));
    unsigned char* scratch = nullptr;
    { int rc = noalloc_scratch(c, &scratch); if (rc) return rc; }
    RTRT_CUDA(c, cudaMemsetAsync(scratch, 0, kNoAllocCounterBytes, c.stream));                      // the array keeps the last call's sums
    RTRT_CUDA(c, cudaMemcpyAsync(scratch + kNoAllocRayOff, h_rays, (size_t)n * sizeof(Ray), cudaMemcpyHostToDevice, c.stream));
    const SceneView sc = make_scene_view(c);
    const uint32_t grid = (n + kNaThreads - 1) / kNaThreads;
    if (c.stackSize <= 16) k_traverse_static<16><<<grid, kNaThreads, 0, c.stream>>>(sc, scratch, n, any_hit);
    else                   k_traverse_static<32><<<grid, kNaThreads, 0, c.stream>>>(sc, scratch, n, any_hit);
    c.launches += 1;
    RTRT_CUDA(c, cudaGetLastError());
    RTRT_CUDA(c, cudaMemcpyAsync(h_hits, scratch + kNoAllocHitOff, (size_t)n * sizeof(Hit), cudaMemcpyDeviceToHost, c.stream));
    unsigned long long cnt[2] = {0, 0};
    RTRT_CUDA(c, cudaMemcpyAsync(cnt, scratch, sizeof(cnt), cudaMemcpyDeviceToHost, c.stream));
    RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    if (node_visits) *node_visits = cnt[0];
    if (leaf_tests) *leaf_tests = cnt[1];
    c.timings.rays_traced = n; c.raysOnDevice = false;
    return RTRT_OK;
}
