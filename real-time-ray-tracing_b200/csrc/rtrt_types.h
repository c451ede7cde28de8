// POD types of the hot path, layout-identical to the reference so that buffers handed out through
// the C ABI can be consumed by (or diffed against) the reference's own structures.
//   Triangle     128 B  src/geometry.h:52-97
//   AABB          24 B  src/geometry.h:14-33   (min then max)
//   AABBCompact   48 B  src/geometry.h:99-152  (two child boxes interleaved in three float4)
//   BVHNode       64 B  src/bvhNode.cuh:5-13
//   Ray           32 B  src/geometry.h:40-50
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define RTRT_HD __host__ __device__ __forceinline__
#else
#define RTRT_HD inline
#endif

namespace rtrt {

struct Float2 { float x, y; };
struct Float3 { float x, y, z; };
struct alignas(16) Float4 { float x, y, z, w; };

struct AABB { Float3 min, max; };

struct alignas(16) Triangle {
    Float3 v1; float w1; Float3 v2; float w2; Float3 v3; float w3; Float3 v4; float w4;
    Float3 n1; float u1; Float3 n2; float u2; Float3 n3; float u3; Float3 n4; float u4;
};

struct alignas(16) BVHNode {
    // AABBCompact
    float box1maxX, box1maxY, box1maxZ, box2minX;
    float box1minX, box1minY, box1minZ, box2minY;
    float box2maxX, box2maxY, box2maxZ, box2minZ;
    // links: leaf child idx = batch-local triangle id (BLAS) or batch id (TLAS)
    uint32_t idxLeft, idxRight, isLeftLeaf, isRightLeaf;
};

struct alignas(16) Ray { Float3 orig; float pad16; Float3 dir; float pad32; };

// one result per traversed ray: the observable outputs of TraverseBvh (src/traverse.h:107-125)
struct Hit { float t, u, v; int32_t tri; };

static_assert(sizeof(Float3) == 12, "Float3");
static_assert(sizeof(AABB) == 24, "AABB");
static_assert(sizeof(Triangle) == 128, "Triangle");
static_assert(sizeof(BVHNode) == 64, "BVHNode");
static_assert(sizeof(Ray) == 32, "Ray");
static_assert(sizeof(Hit) == 16, "Hit");

constexpr uint32_t kBatchSize = 1024;      // src/kernel.cuh:579  BatchSize
constexpr float kRayMax = 10e10f;          // src/kernel.cuh:69   RayMax

}  // namespace rtrt
