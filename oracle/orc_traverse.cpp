// ORACLE — TEST INFRASTRUCTURE ONLY (parity pinned against oracle/_ref; see orc_types.h header).
// C entry point for batched ray traversal: one TraverseBvh (src/traverse.h:107-253) per ray with the
// state initialisation of RaySceneIntersect (src/traverse.cuh:86-101).
#include "orc_traverse.h"
#include <thread>
#include <vector>

using namespace orc;

extern "C" {

// rays: n x {orig.xyz, pad, dir.xyz, pad} (src/geometry.h:40-50); hits: n x {t,u,v,tri}
// counters (optional, may be NULL): [0] += node visits, [1] += leaf tests
int orc_traverse(const void* triangles, const void* nodes, const void* tlas, u32 numBatches,
                 const void* rays, u32 n, void* hits, int anyHit, int stackSize, int maxLoop,
                 unsigned long long* counters, int threads) {
    Scene sc{(const Triangle*)triangles, (const BVHNode*)nodes, (const BVHNode*)tlas, numBatches, stackSize, maxLoop};
    const Ray* R = (const Ray*)rays; Hit* H = (Hit*)hits;
    if (threads < 1) threads = 1;
    std::vector<unsigned long long> nv(threads, 0), lt(threads, 0);
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
        pool.emplace_back([&, t]() {
            u32 lo = (u32)((unsigned long long)n * t / threads), hi = (u32)((unsigned long long)n * (t + 1) / threads);
            for (u32 i = lo; i < hi; ++i) {
                Ray ray = R[i];
                F3 inv = SafeDivide3f(f3(1.0f), ray.dir);          // src/traverse.cuh:88
                HitRecord r = init_hit();
                TraverseBvh(sc, ray, inv, r, anyHit);
                H[i] = Hit{r.t, r.uv.x, r.uv.y, r.objectIdx};
                nv[t] += r.nodeVisits; lt[t] += r.leafTests;
            }
        });
    for (auto& th : pool) th.join();
    if (counters) for (int t = 0; t < threads; ++t) { counters[0] += nv[t]; counters[1] += lt[t]; }
    return 0;
}

}  // extern "C"

// ---- small probes used by the CPU tests to pin the oracle itself ----
namespace orc { u32 MortonCode3D(u32 x, u32 y, u32 z); }

extern "C" {

u32 orc_morton30(u32 x, u32 y, u32 z) { return orc::MortonCode3D(x, y, z); }
u32 orc_cuda_f2u(float f) { return orc::cuda_f2u(f); }

// closest hit by testing EVERY triangle with the same watertight routine (no hierarchy involved)
int orc_brute_force_closest(const void* triangles, u32 numTris, const void* rays, u32 n, void* hits, int threads) {
    const Triangle* T = (const Triangle*)triangles; const Ray* R = (const Ray*)rays; Hit* H = (Hit*)hits;
    if (threads < 1) threads = 1;
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
        pool.emplace_back([&, t]() {
            for (u32 i = t; i < n; i += threads) {
                float best = RayMax, u = 0, v = 0, e = 0; int tri = -1;
                for (u32 k = 0; k < numTris; ++k) {
                    float tt;
                    if (RayTriangleWatertight(R[i], T[k], best, tt, u, v, e) && tt < best) { best = tt; tri = (int)k; }
                }
                H[i] = Hit{best, u, v, tri};
            }
        });
    for (auto& th : pool) th.join();
    return 0;
}

}  // extern "C"
