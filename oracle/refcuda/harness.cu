// oracle/refcuda/harness.cu — extern "C" launchers around the REFERENCE'S OWN kernels (test / baseline infrastructure).
// The kernel bodies come from /root/reference/src/*.cuh (never copied into this repo); this file only reproduces the launch
// sequence of src/bvh.cu:7-97 and a one-thread-per-ray wrapper around TraverseBvh (src/traverse.h:107-253) with the
// set-up of RaySceneIntersect (src/traverse.cuh:86-101).  Built by oracle/refcuda/build_ref.py into oracle/_ref/.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#define RayMax 10e10f            /* src/kernel.cuh:69 (kernel.cuh itself drags in the Vulkan/GLFW-era includes) */
#include "cudaError.cuh"
#include "linearMath.h"
#include "geometry.h"
#include "bvhNode.cuh"
#include "updateGeometry.cuh"
#include "radixSort.cuh"
#include "buildBVH.cuh"
#include "traverse.h"

static_assert(sizeof(Float3) == 12 && sizeof(Float4) == 16 && sizeof(AABB) == 24 && sizeof(Triangle) == 128 && sizeof(BVHNode) == 64,
              "type shim changed a reference layout");

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "refcuda: %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return -1; } } while (0)

constexpr uint KernelSize = 256, KernalBatchSize = 4, BatchSize = 1024;     // src/kernel.cuh:579-581

struct RefRay { float orig[3]; float pad16; float dir[3]; float pad32; };
struct RefHit { float t, u, v; int tri; };

__global__ void RefTraverse(Triangle* triangles, BVHNode* bvhNodes, BVHNode* tlasBvhNodes, uint batchSize, const RefRay* rays, uint n, RefHit* hits,
                            int bvhNodesSize, int tlasSize, int trianglesSize) {
    uint i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Ray ray;
    ray.orig = Float3(rays[i].orig[0], rays[i].orig[1], rays[i].orig[2]);
    ray.dir = Float3(rays[i].dir[0], rays[i].dir[1], rays[i].dir[2]);
    Float3 invRayDir = SafeDivide3f(Float3(1.0f), ray.dir);                 // src/traverse.cuh:88
    float errorP = 1e-7f, errorT = 1e-7f, t = RayMax, rayOffset = 1e-7f;    // src/traverse.cuh:86,100-101
    Float2 uv(0.0f, 0.0f);
    int objectIdx = -1;
    Float3 n0(0.0f), fakeN(0.0f), p(RayMax);
    TraverseBvh(triangles, bvhNodes, tlasBvhNodes, batchSize, ray, invRayDir, errorP, errorT, t, uv, objectIdx, n0, fakeN, p, rayOffset,
                bvhNodesSize, tlasSize, trianglesSize);
    hits[i].t = t; hits[i].u = uv.x; hits[i].v = uv.y; hits[i].tri = objectIdx;
}

extern "C" {

// Device pointers in, reference kernels run on the default stream; ms_out[0] = build time by CUDA events.
// Arrays: sizes as in src/init.cu:139-187 (morton / reorderIdx hold batchCount*1024 entries, pre-filled here with 0xFF like
// src/kernel.cu:278-280).  Reference limits apply: batchCount < 1024.
int refcuda_build_bvh(Float3* vertexBuffer, uint* indexBuffer, Float3* normalBuffer, uint triCount,
                      Triangle* triangles, AABB* aabbs, uint* morton, uint* reorderIdx, BVHNode* bvhNodes,
                      AABB* tlasAabbs, uint* tlasMorton, uint* tlasReorderIdx, BVHNode* tlasBvhNodes,
                      uint* triCountArray, uint* batchCountArray, int iterations, float* ms_out) {
    uint triCountPadded = (triCount + 3u) / 4u * 4u;
    uint triCountPadded2 = (triCount + BatchSize - 1) / BatchSize * BatchSize;
    uint batchCount = triCountPadded2 / BatchSize;
    if (batchCount >= BatchSize) return -2;                                  // assert(batchCount < BatchSize), src/init.cu:126
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float total = 0.0f;
    for (int it = 0; it < iterations; ++it) {
        CK(cudaMemset(morton, 0xFF, (size_t)triCountPadded2 * sizeof(uint)));       // src/kernel.cu:279 (whole padded range here)
        CK(cudaMemset(tlasMorton, 0xFF, BatchSize * sizeof(uint)));                  // src/kernel.cu:280
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0));
        UpdateSceneGeometry<KernelSize, KernalBatchSize><<<batchCount, KernelSize>>>(vertexBuffer, indexBuffer, normalBuffer, triangles, aabbs, morton, triCountArray, 0.0f);
        RadixSort<KernelSize, KernalBatchSize><<<batchCount, KernelSize>>>(morton, reorderIdx);
        BuildLBVH<KernelSize, KernalBatchSize><<<batchCount, KernelSize>>>(bvhNodes, aabbs, morton, reorderIdx, triCountArray);
        UpdateTLAS<KernelSize, KernalBatchSize, BatchSize><<<1, KernelSize>>>(bvhNodes, tlasAabbs, tlasMorton, batchCountArray, triCountPadded);
        RadixSort<KernelSize, KernalBatchSize><<<1, KernelSize>>>(tlasMorton, tlasReorderIdx);
        BuildLBVH<KernelSize, KernalBatchSize><<<1, KernelSize>>>(tlasBvhNodes, tlasAabbs, tlasMorton, tlasReorderIdx, batchCountArray);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaGetLastError());
        float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (it > 0 || iterations == 1) total += ms;
    }
    if (ms_out) ms_out[0] = total / (iterations > 1 ? iterations - 1 : 1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return 0;
}

int refcuda_traverse(Triangle* triangles, BVHNode* bvhNodes, BVHNode* tlasBvhNodes, uint batchCount, const RefRay* rays, uint n, RefHit* hits,
                     int iterations, float* ms_out) {
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float total = 0.0f;
    for (int it = 0; it < iterations; ++it) {
        CK(cudaEventRecord(e0));
        RefTraverse<<<(n + 63) / 64, 64>>>(triangles, bvhNodes, tlasBvhNodes, BatchSize, rays, n, hits, (int)(batchCount * BatchSize), (int)batchCount,
                                           (int)(batchCount * BatchSize));
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaGetLastError());
        float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (it > 0 || iterations == 1) total += ms;
    }
    if (ms_out) ms_out[0] = total / (iterations > 1 ? iterations - 1 : 1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return 0;
}

}  // extern "C"
