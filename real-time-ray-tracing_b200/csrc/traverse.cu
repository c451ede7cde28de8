// Stage 2 (traversal-only entry point): one TraverseBvh per ray — rtrt_traverse / rtrt_traverse_host.
// Replaces the TraverseBvh call inside RaySceneIntersect (src/traverse.cuh:64-121) as a stand-alone
// query kernel so per-ray (t, tri, u, v) can be diffed against the oracle.
#include "wide_bvh.cuh"
#include "context.h"

namespace rtrt {

constexpr int kTravThreads = 128;

template <int STACK, bool COUNT>
__global__ void __launch_bounds__(kTravThreads) k_traverse(SceneView sc, const Ray* __restrict__ rays, uint32_t n,
                                                           Hit* __restrict__ hits, int anyHit, unsigned long long* counters) {
    __shared__ uint32_t sLink[STACK * kTravThreads];
    __shared__ float sT[STACK * kTravThreads];
    const uint32_t i = blockIdx.x * kTravThreads + threadIdx.x;
    if (i >= n) return;
    // scene box = merged box of TLAS node 0 (src/traverse.h:127-129)
    const float4 a = __ldg(sc.tlasRows), b = __ldg(sc.tlasRows + 1), c = __ldg(sc.tlasRows + 2);
    AABB scene;
    scene.max = Float3{sel_max(a.x, c.x), sel_max(a.y, c.y), sel_max(a.z, c.z)};
    scene.min = Float3{sel_min(b.x, a.w), sel_min(b.y, b.w), sel_min(b.z, c.w)};
    const float4* r4 = reinterpret_cast<const float4*>(rays + i);
    const float4 o = __ldg(r4), d = __ldg(r4 + 1);
    const RayPrep rp = prepare_ray(Float3{o.x, o.y, o.z}, Float3{d.x, d.y, d.z}, scene);
    const TraverseResult res = traverse_bvh<STACK, COUNT>(sc, rp, anyHit, sLink + threadIdx.x, sT + threadIdx.x, kTravThreads);
    reinterpret_cast<float4*>(hits)[i] = make_float4(res.t, res.u, res.v, __int_as_float(res.tri));
    if (COUNT) {
        atomicAdd(counters, (unsigned long long)res.nodeVisits);
        atomicAdd(counters + 1, (unsigned long long)res.leafTests);
    }
}

// Reference layout with a 64-entry stack in local memory: the uncapped checker of the tests (stack_size 33..64 together with a
// large max_loop): no push is ever dropped on the scenes tested, so this is "the reference's search without its two caps".
template <bool COUNT>
__global__ void __launch_bounds__(kTravThreads) k_traverse_deep(SceneView sc, const Ray* __restrict__ rays, uint32_t n,
                                                                Hit* __restrict__ hits, int anyHit, unsigned long long* counters) {
    const uint32_t i = blockIdx.x * kTravThreads + threadIdx.x;
    if (i >= n) return;
    uint32_t lLink[64]; float lT[64];
    const float4 a = __ldg(sc.tlasRows), b = __ldg(sc.tlasRows + 1), c = __ldg(sc.tlasRows + 2);
    AABB scene;
    scene.max = Float3{sel_max(a.x, c.x), sel_max(a.y, c.y), sel_max(a.z, c.z)};
    scene.min = Float3{sel_min(b.x, a.w), sel_min(b.y, b.w), sel_min(b.z, c.w)};
    const float4* r4 = reinterpret_cast<const float4*>(rays + i);
    const float4 o = __ldg(r4), d = __ldg(r4 + 1);
    const RayPrep rp = prepare_ray(Float3{o.x, o.y, o.z}, Float3{d.x, d.y, d.z}, scene);
    const TraverseResult res = traverse_bvh<64, COUNT>(sc, rp, anyHit, lLink, lT, 1);
    reinterpret_cast<float4*>(hits)[i] = make_float4(res.t, res.u, res.v, __int_as_float(res.tri));
    if (COUNT) { atomicAdd(counters, (unsigned long long)res.nodeVisits); atomicAdd(counters + 1, (unsigned long long)res.leafTests); }
}

// Wide layout (option wide_bvh): closest hit over the 8-wide compressed nodes; counters[2] += stack overflows (must stay 0).
template <bool COUNT>
__global__ void __launch_bounds__(kTravThreads) k_traverse_wide(SceneView sc, const Ray* __restrict__ rays, uint32_t n,
                                                                Hit* __restrict__ hits, int anyHit, unsigned long long* counters) {
    __shared__ uint2 sRaw[16 * kTravThreads];
    const uint32_t i = blockIdx.x * kTravThreads + threadIdx.x;
    WideTrav tr;
    uint2 spill[WideTrav::kSpill];
    tr.setup(sc, sRaw, threadIdx.x, kTravThreads, spill);
    bool active = i < n;                                           // whole warps step together (wide_step votes)
    if (active) {
        const float4* r4 = reinterpret_cast<const float4*>(rays + i);
        const float4 o = __ldg(r4), d = __ldg(r4 + 1);
        tr.begin(sc, Float3{o.x, o.y, o.z}, Float3{d.x, d.y, d.z});
    }
    while (__any_sync(0xffffffffu, active)) {
        if (tr.template step_warp<COUNT>(sc, anyHit, active)) active = false;
    }
    if (i >= n) return;
    reinterpret_cast<float4*>(hits)[i] = make_float4(tr.t(), tr.u(), tr.v(), __int_as_float(tr.tri()));
    if (COUNT) {
        atomicAdd(counters, (unsigned long long)tr.nodeVisits()); atomicAdd(counters + 1, (unsigned long long)tr.leafTests());
        if (tr.s.overflow) atomicAdd(counters + 2, (unsigned long long)tr.s.overflow);
    }
}

int ensure_wide(Context& c) {
    if (c.wideBvh && !c.wideValid) return launch_wide_build(c);
    return RTRT_OK;
}

SceneView make_scene_view(const Context& c) {
    SceneView sc;
    sc.triRows = reinterpret_cast<const float4*>(c.dTriangles);
    sc.nodeRows = reinterpret_cast<const float4*>(c.dNodes);
    sc.tlasRows = reinterpret_cast<const float4*>(c.dTlasNodes);
    sc.numBatches = c.numBatches;
    sc.stackSize = c.stackSize;
    sc.maxLoop = c.maxLoop;
    sc.wide.nodes = reinterpret_cast<const float4*>(c.dWideNodes); sc.wide.tris = reinterpret_cast<const float4*>(c.dWideTris); sc.wide.root = c.wideRoot;
    return sc;
}

int launch_traverse(Context& c, const Ray* dRays, uint32_t n, Hit* dHits, int anyHit, unsigned long long* dCounters) {
    if (!c.bvhValid) return c.fail(RTRT_ERR_STATE, "rtrt_traverse: BVH not built");
    if (n == 0) return RTRT_OK;
    { int rc = ensure_wide(c); if (rc) return rc; }
    if (c.wideBvh) { int rc = wide_join(c); if (rc) return rc; }
    const SceneView sc = make_scene_view(c);
    const uint32_t grid = (n + kTravThreads - 1) / kTravThreads;
    if (c.wideBvh) {
        if (dCounters) k_traverse_wide<true><<<grid, kTravThreads, 0, c.stream>>>(sc, dRays, n, dHits, anyHit, dCounters);
        else           k_traverse_wide<false><<<grid, kTravThreads, 0, c.stream>>>(sc, dRays, n, dHits, anyHit, nullptr);
    } else if (c.stackSize > 32) {
        if (dCounters) k_traverse_deep<true><<<grid, kTravThreads, 0, c.stream>>>(sc, dRays, n, dHits, anyHit, dCounters);
        else           k_traverse_deep<false><<<grid, kTravThreads, 0, c.stream>>>(sc, dRays, n, dHits, anyHit, nullptr);
    } else if (dCounters) {
        if (c.stackSize <= 16) k_traverse<16, true><<<grid, kTravThreads, 0, c.stream>>>(sc, dRays, n, dHits, anyHit, dCounters);
        else                   k_traverse<32, true><<<grid, kTravThreads, 0, c.stream>>>(sc, dRays, n, dHits, anyHit, dCounters);
    } else {
        if (c.stackSize <= 16) k_traverse<16, false><<<grid, kTravThreads, 0, c.stream>>>(sc, dRays, n, dHits, anyHit, nullptr);
        else                   k_traverse<32, false><<<grid, kTravThreads, 0, c.stream>>>(sc, dRays, n, dHits, anyHit, nullptr);
    }
    c.launches += 1;
    RTRT_CUDA(c, cudaGetLastError());
    return RTRT_OK;
}

}  // namespace rtrt
