// ORACLE — TEST INFRASTRUCTURE ONLY.  Parity status: PINNED against the reference's own CUDA kernels compiled for sm_100
// (oracle/refcuda -> oracle/_ref/librefcuda.so): build + traversal bit-exact, sky / PathTrace / denoise within the tolerances
// of DESIGN.md §2 (tests/test_refcuda_gpu.py, tests/test_refcuda_frame_gpu.py).  The reference ships no golden vectors, KATs or
// CPU path for this hot path.
//
// CPU restatement of the reference's hot-path arithmetic (wangkepfe/Real-Time-Ray-Tracing).
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// link or load this code.  The product (real-time-ray-tracing_b200/) never does.
//
// Build: g++ -O2 -ffp-contract=off -fno-fast-math (so that a*b+c is never fused; fmaf is used
// only where the reference calls FMA explicitly: src/linearMath.h:41-98).
#pragma once
#include <cstdint>
#include <cstring>
#include <cmath>
#include <cfloat>

namespace orc {

typedef uint32_t u32;

// ---- src/linearMath.h:249-311 (Float3), 100-130 (Float2) : plain layout-identical PODs ----
struct F2 { float x, y; };
struct F3 {
    float x, y, z;
    float  operator[](int i) const { return (&x)[i]; }
    float& operator[](int i)       { return (&x)[i]; }
};
struct F4 { float x, y, z, w; };

static inline F3 f3(float a) { return F3{a, a, a}; }
static inline F3 f3(float a, float b, float c) { return F3{a, b, c}; }
static inline F3 operator+(F3 a, F3 b) { return F3{a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline F3 operator-(F3 a, F3 b) { return F3{a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline F3 operator*(F3 a, F3 b) { return F3{a.x * b.x, a.y * b.y, a.z * b.z}; }
static inline F3 operator/(F3 a, F3 b) { return F3{a.x / b.x, a.y / b.y, a.z / b.z}; }
static inline F3 operator*(F3 a, float s) { return F3{a.x * s, a.y * s, a.z * s}; }
static inline F3 operator*(float s, F3 a) { return F3{a.x * s, a.y * s, a.z * s}; }
static inline F3 operator/(F3 a, float s) { return F3{a.x / s, a.y / s, a.z / s}; }
static inline F3 operator+(F3 a, float s) { return F3{a.x + s, a.y + s, a.z + s}; }
static inline F3 operator-(F3 a, float s) { return F3{a.x - s, a.y - s, a.z - s}; }
static inline F3 operator-(F3 a) { return F3{-a.x, -a.y, -a.z}; }
static inline F2 operator+(F2 a, F2 b) { return F2{a.x + b.x, a.y + b.y}; }
static inline F2 operator-(F2 a, F2 b) { return F2{a.x - b.x, a.y - b.y}; }
static inline F2 operator*(F2 a, F2 b) { return F2{a.x * b.x, a.y * b.y}; }
static inline F2 operator/(F2 a, F2 b) { return F2{a.x / b.x, a.y / b.y}; }
static inline F2 operator*(F2 a, float s) { return F2{a.x * s, a.y * s}; }
static inline F2 operator+(F2 a, float s) { return F2{a.x + s, a.y + s}; }
static inline F2 operator-(F2 a, float s) { return F2{a.x - s, a.y - s}; }
static inline F2 operator/(F2 a, float s) { return F2{a.x / s, a.y / s}; }

// The reference's scalar min(a, b) / max(a, b): for two floats overload resolution picks CUDA's non-template ::min / ::max (= fminf /
// fmaxf, one FMNMX on the device) over the select-style templates of src/linearMath.h:25-31, so: NaN loses and -0 < +0.  (glibc's
// fminf is not sign-of-zero aware, hence written out.)  Pinned by tests/test_refcuda_gpu.py::test_signed_zero_boxes_bitwise.
static inline float rmax(float a, float b) { if (a != a) return b; if (b != b) return a; if (a == b) return std::signbit(a) ? b : a; return a > b ? a : b; }
static inline float rmin(float a, float b) { if (a != a) return b; if (b != b) return a; if (a == b) return std::signbit(a) ? a : b; return a < b ? a : b; }
static inline float rmax3(float a, float b, float c) { return rmax(rmax(a, b), c); }
static inline float rmin3(float a, float b, float c) { return rmin(rmin(a, b), c); }
static inline F3 min3f(F3 a, F3 b) { return F3{rmin(a.x, b.x), rmin(a.y, b.y), rmin(a.z, b.z)}; }   // linearMath.h:465
static inline F3 max3f(F3 a, F3 b) { return F3{rmax(a.x, b.x), rmax(a.y, b.y), rmax(a.z, b.z)}; }   // linearMath.h:466
static inline F3 abs3(F3 v) { return F3{fabsf(v.x), fabsf(v.y), fabsf(v.z)}; }                      // linearMath.h:459
static inline float getmax(F3 v) { return rmax(rmax(v.x, v.y), v.z); }                              // linearMath.h:305

// src/linearMath.h:41-53 — FMA + difference of products
static inline float FMA(float a, float b, float c) { return fmaf(a, b, c); }
static inline float dop(float a, float b, float c, float d) {
    float cd = c * d;
    float err = FMA(-c, d, cd);
    float r = FMA(a, b, -cd);
    return r + err;
}
// src/linearMath.h:64-91 — compensated inner product (TwoProd / TwoSum)
struct CF { float v, err; };
static inline CF TwoProd(float a, float b) { float ab = a * b; return CF{ab, FMA(a, b, -ab)}; }
static inline CF TwoSum(float a, float b) {
    float s = a + b, delta = s - a;
    return CF{s, (a - (s - delta)) + (b - delta)};
}
// dot(Float3,Float3) = (float)InnerProduct(x,x', y,y', z,z')   src/linearMath.h:484
static inline float dot(F3 a, F3 b) {
    CF zz = TwoProd(a.z, b.z);                      // InnerProduct(z,z')
    CF yy = TwoProd(a.y, b.y);
    CF s1 = TwoSum(yy.v, zz.v);
    CF t1 = CF{s1.v, yy.err + (zz.err + s1.err)};   // InnerProduct(y,y',z,z')
    CF xx = TwoProd(a.x, b.x);
    CF s0 = TwoSum(xx.v, t1.v);
    CF t0 = CF{s0.v, xx.err + (t1.err + s0.err)};
    return t0.v + t0.err;
}
static inline float dot2(F2 a, F2 b) { return a.x * b.x + a.y * b.y; }                              // linearMath.h:482
// cross = three dop()s   src/linearMath.h:468
static inline F3 cross(F3 a, F3 b) {
    return F3{dop(a.y, b.z, a.z, b.y), dop(a.z, b.x, a.x, b.z), dop(a.x, b.y, a.y, b.x)};
}
static inline F3 normalize(F3 v) {                                                                  // linearMath.h:461
    float n = sqrtf(v.x * v.x + v.y * v.y + v.z * v.z);
    return F3{v.x / n, v.y / n, v.z / n};
}
static inline float length2(F3 v) { return v.x * v.x + v.y * v.y + v.z * v.z; }
static inline float length(F3 v) { return sqrtf(v.x * v.x + v.y * v.y + v.z * v.z); }
static inline float clampf(float a, float lo = 0.0f, float hi = 1.0f) { return a < lo ? lo : a > hi ? hi : a; }
// src/linearMath.h:721-723
static inline float SafeDivide(float a, float b) {
    float eps = 1e-20f;
    return a / ((fabsf(b) > eps) ? b : copysignf(eps, b));
}
static inline F3 SafeDivide3f(F3 a, F3 b) { return F3{SafeDivide(a.x, b.x), SafeDivide(a.y, b.y), SafeDivide(a.z, b.z)}; }

// ---- src/precision.cuh:18-70 ----
static inline float MachineEpsilon() { return 1.1920928955078125e-07f; }              // 2^-23
static inline float ErrGamma(int n) { return (n * MachineEpsilon()) / (1.0f - n * MachineEpsilon()); }
static inline float TwoPowerMinus23() { return 1.1920928955078125e-07f; }
static inline float TwoPowerMinus24() { return 5.9604644775390625e-08f; }
static inline float pP() { return 1.0f + TwoPowerMinus23(); }
static inline float pM() { return 1.0f - TwoPowerMinus23(); }
static inline float up(float a) { return a > 0.0f ? a * pP() : a * pM(); }
static inline float dn(float a) { return a > 0.0f ? a * pM() : a * pP(); }
static inline float Up(float a) { return a * pP(); }
static inline float Dn(float a) { return a * pM(); }
static inline F3 Up(F3 a) { return F3{Up(a.x), Up(a.y), Up(a.z)}; }
static inline F3 Dn(F3 a) { return F3{Dn(a.x), Dn(a.y), Dn(a.z)}; }

// CUDA float->uint conversion semantics (cvt.rzi.u32.f32): NaN -> 0, negative -> 0, saturating.
// The reference relies on this at src/updateGeometry.cuh:256-258 (0/0 on a flat axis).
static inline u32 cuda_f2u(float f) {
    if (!(f == f)) return 0u;
    if (f <= 0.0f) return 0u;
    if (f >= 4294967296.0f) return 0xFFFFFFFFu;
    return (u32)f;
}

// ---- src/geometry.h:14-33 AABB: min then max, 24 bytes ----
struct AABB {
    F3 min, max;
    float operator[](int i) const { return (&min.x)[i]; }
};
static inline AABB aabb_empty() { return AABB{f3(FLT_MAX), f3(-FLT_MAX)}; }

// ---- src/geometry.h:52-97 Triangle: 128 bytes ----
struct Triangle {
    F3 v1; float w1; F3 v2; float w2; F3 v3; float w3; F3 v4; float w4;
    F3 n1; float u1; F3 n2; float u2; F3 n3; float u3; F3 n4; float u4;
};
// ---- src/geometry.h:99-152 AABBCompact (48 B) + src/bvhNode.cuh:5-13 BVHNode (64 B) ----
struct BVHNode {
    float box1maxX, box1maxY, box1maxZ, box2minX;
    float box1minX, box1minY, box1minZ, box2minY;
    float box2maxX, box2maxY, box2maxZ, box2minZ;
    u32 idxLeft, idxRight, isLeftLeaf, isRightLeaf;
};
static_assert(sizeof(AABB) == 24, "AABB layout");
static_assert(sizeof(Triangle) == 128, "Triangle layout");
static_assert(sizeof(BVHNode) == 64, "BVHNode layout");

static inline AABB node_left(const BVHNode& n)  { return AABB{F3{n.box1minX, n.box1minY, n.box1minZ}, F3{n.box1maxX, n.box1maxY, n.box1maxZ}}; }
static inline AABB node_right(const BVHNode& n) { return AABB{F3{n.box2minX, n.box2minY, n.box2minZ}, F3{n.box2maxX, n.box2maxY, n.box2maxZ}}; }
static inline void node_set_left(BVHNode& n, const AABB& b)  { n.box1maxX = b.max.x; n.box1maxY = b.max.y; n.box1maxZ = b.max.z; n.box1minX = b.min.x; n.box1minY = b.min.y; n.box1minZ = b.min.z; }
static inline void node_set_right(BVHNode& n, const AABB& b) { n.box2maxX = b.max.x; n.box2maxY = b.max.y; n.box2maxZ = b.max.z; n.box2minX = b.min.x; n.box2minY = b.min.y; n.box2minZ = b.min.z; }
// AABBCompact::GetMerged  src/geometry.h:107-110
static inline AABB node_merged(const BVHNode& n) {
    return AABB{F3{rmin(n.box1minX, n.box2minX), rmin(n.box1minY, n.box2minY), rmin(n.box1minZ, n.box2minZ)},
                F3{rmax(n.box1maxX, n.box2maxX), rmax(n.box1maxY, n.box2maxY), rmax(n.box1maxZ, n.box2maxZ)}};
}

struct Ray { F3 orig; float pad0; F3 dir; float pad1; };   // src/geometry.h:40-50

// one result per ray, the observation point of TraverseBvh  (src/traverse.h:107-125)
struct Hit {
    float t;          // RayMax (1e11) on miss
    float u, v;
    int   tri;        // objectIdx = blasOffset*1024 + batch-local id ; -1 on miss
};

static const float RayMax = 10e10f;   // src/kernel.cuh:69
static const u32 kBatch = 1024;       // src/kernel.cuh:579-581 BatchSize

}  // namespace orc
