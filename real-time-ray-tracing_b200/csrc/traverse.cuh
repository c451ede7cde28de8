// Stage 2 core — two-level BVH traversal for sm_100a (no RT cores on B200: everything is SIMT).
//
// Semantics restated from (results must match bit for bit: closest t, triangle id, u, v):
//   src/traverse.h:107-253    TraverseBvh      ordered DFS, near child first, far child pushed with tNear,
//                                              pop-time culling (TestForFinish :88-105), 16-entry stack with
//                                              dropped pushes (:32-40), 1024-iteration cap (:132)
//   src/geometry.cuh:519-616  conservative slab test with pre-rounded origins / reciprocal directions
//   src/geometry.cuh:375-472  watertight ray/triangle test with double-precision fallback
// What is different from the reference (none of it changes a result bit):
//   * the per-ray box helper is kept per world axis (sign-selected near/far plane) instead of the
//     kx/ky/kz-permuted form — max/min over three axes do not depend on the order;
//   * nodes are fetched as 4 x 16-byte read-only loads, leaf tests fetch only the three position rows
//     (48 of the 128 bytes); normals are read once, for the accepted hit, after the loop;
//   * stack entries are one packed link word + tNear.
// Compiled with --fmad=false: a*b+c must not be contracted (the oracle uses -ffp-contract=off).
#pragma once
#include "rtrt_types.h"
#include "bvh_block.cuh"   // sel_min / sel_max
#include <cuda_runtime.h>
#include <float.h>

namespace rtrt {

// the 8-wide compressed layout (wide_bvh.cuh), null when option wide_bvh is off
struct WideView {
    const float4* __restrict__ nodes;            // 8 float4 rows (128 B) per node, 80 B used
    const float4* __restrict__ tris;             // 3 float4 rows per triangle, row0.w = original triangle index
    uint32_t root;                               // index of the top-level root node
};

struct SceneView {
    const float4* __restrict__ triRows;     // Triangle[] viewed as 8 float4 rows per triangle
    const float4* __restrict__ nodeRows;    // BVHNode[]  viewed as 4 float4 rows per node (BLAS, 1024 per batch)
    const float4* __restrict__ tlasRows;    // TLAS nodes
    uint32_t numBatches;
    int stackSize, maxLoop;
    WideView wide;
};

__device__ __forceinline__ float comp(const Float3& v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : v.z); }
__device__ __forceinline__ float xor_sign(float a, uint32_t s) { return __uint_as_float(__float_as_uint(a) ^ s); }

// src/precision.cuh:44-66
__device__ __forceinline__ float f_p() { return 1.0f + 1.1920928955078125e-07f; }
__device__ __forceinline__ float f_m() { return 1.0f - 1.1920928955078125e-07f; }
__device__ __forceinline__ float rnd_up(float a) { return a > 0.0f ? a * f_p() : a * f_m(); }   // up()
__device__ __forceinline__ float rnd_dn(float a) { return a > 0.0f ? a * f_m() : a * f_p(); }   // dn()
__device__ __forceinline__ float Upf(float a) { return a * f_p(); }
__device__ __forceinline__ float Dnf(float a) { return a * f_m(); }

__device__ __forceinline__ int max_dim3(float ax, float ay, float az) {     // src/geometry.cuh:375-387
    if (ax > ay) { return az > ax ? 2 : 0; }
    return az > ay ? 2 : 1;
}

// src/linearMath.h:721-723
__device__ __forceinline__ float safe_divide(float a, float b) {
    const float eps = 1e-20f;
    return a / ((fabsf(b) > eps) ? b : copysignf(eps, b));
}

// Per-ray constants of both tests.
struct RayPrep {
    Float3 org, dir;
    // box test, per world axis a: tNear_a = (box[nearIsMax ? max : min] - orgNear) * rdirNear  (and far likewise)
    Float3 orgNear, orgFar, rdirNear, rdirFar;
    bool negX, negY, negZ;
    // watertight test
    int kx, ky, kz;
    float Sx, Sy, Sz;
};

__device__ __forceinline__ void set_comp(Float3& v, int k, float f) { if (k == 0) v.x = f; else if (k == 1) v.y = f; else v.z = f; }

// CreateRayBoxIntersectionHelper (src/geometry.cuh:519-583) + the ray-only part of RayTriangleWatertight (:411-421)
__device__ __forceinline__ RayPrep prepare_ray(const Float3& org, const Float3& dir, const AABB& sceneBox) {
    RayPrep r;
    r.org = org; r.dir = dir;
    const Float3 inv{safe_divide(1.0f, dir.x), safe_divide(1.0f, dir.y), safe_divide(1.0f, dir.z)};   // src/traverse.cuh:88
    int kz = max_dim3(fabsf(dir.x), fabsf(dir.y), fabsf(dir.z));
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    // The reference swaps kx and ky when dir[kz] < 0 (src/geometry.cuh:415-416) to keep the winding.  The swap negates
    // U, V, W, det and T exactly (fl(a-b) = -fl(b-a), no FMA contraction in this TU), so the sign tests, u = U/det,
    // v = V/det and t = T/det are bit-identical without it; leaving it out makes (kx, ky, kz) a pure rotation of
    // (x, y, z), which ray_triangle selects with predicated moves instead of divergent branches.
    r.kx = kx; r.ky = ky; r.kz = kz;
    r.Sx = comp(dir, kx) / comp(dir, kz);
    r.Sy = comp(dir, ky) / comp(dir, kz);
    r.Sz = 1.0f / comp(dir, kz);
    r.negX = dir.x < 0.0f; r.negY = dir.y < 0.0f; r.negZ = dir.z < 0.0f;

    const float eps = 5.0f * 5.9604644775390625e-08f;          // 5 * 2^-24
    const Float3 lower{Dnf(fabsf(org.x - sceneBox.min.x)), Dnf(fabsf(org.y - sceneBox.min.y)), Dnf(fabsf(org.z - sceneBox.min.z))};
    const Float3 upper{Upf(fabsf(org.x - sceneBox.max.x)), Upf(fabsf(org.y - sceneBox.max.y)), Upf(fabsf(org.z - sceneBox.max.z))};
    const float max_z = sel_max(comp(lower, kz), comp(upper, kz));
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float o = comp(org, a), d = comp(dir, a);
        float on, of;
        if (a == kz) { on = o; of = o; }
        else {
            const float err_near = Upf(comp(lower, a) + max_z);
            const float err_far = Upf(comp(upper, a) + max_z);
            on = rnd_up(o + Upf(eps * err_near));
            of = rnd_dn(o - Upf(eps * err_far));
            if (d < 0.0f) { const float t = on; on = of; of = t; }
        }
        set_comp(r.orgNear, a, on); set_comp(r.orgFar, a, of);
        set_comp(r.rdirNear, a, Dnf(Dnf(comp(inv, a))));
        set_comp(r.rdirFar, a, Upf(Upf(comp(inv, a))));
    }
    return r;
}

// RayAABBIntersect (src/geometry.cuh:585-603) on explicit min/max
__device__ __forceinline__ bool ray_box(const RayPrep& r, float mnx, float mny, float mnz, float mxx, float mxy, float mxz, float& tNear) {
    const float tnx = ((r.negX ? mxx : mnx) - r.orgNear.x) * r.rdirNear.x;
    const float tny = ((r.negY ? mxy : mny) - r.orgNear.y) * r.rdirNear.y;
    const float tnz = ((r.negZ ? mxz : mnz) - r.orgNear.z) * r.rdirNear.z;
    const float tfx = ((r.negX ? mnx : mxx) - r.orgFar.x) * r.rdirFar.x;
    const float tfy = ((r.negY ? mny : mxy) - r.orgFar.y) * r.rdirFar.y;
    const float tfz = ((r.negZ ? mnz : mxz) - r.orgFar.z) * r.rdirFar.z;
    // fmaxf / fminf like the reference: its scalar max() / min() calls resolve to CUDA's overloads (bvh_block.cuh), one FMNMX each
    const float tn = fmaxf(fmaxf(tnx, tny), tnz);
    const float tf = fminf(fminf(tfx, tfy), tfz);
    const bool hit = tn <= tf && (tf > 0.0f);
    tNear = fmaxf(tn, 0.0f);
    return hit;
}

// branch-free select (the compiler turns the ternary chains of comp() into divergent branches)
__device__ __forceinline__ float fsel(int p, float a, float b) {
    float r;
    asm("{ .reg .pred q; setp.ne.s32 q, %1, 0; selp.f32 %0, %2, %3, q; }" : "=f"(r) : "r"(p), "f"(a), "f"(b));
    return r;
}

// RayTriangleWatertight (src/geometry.cuh:406-472); v1,v2,v3 = triangle rows 0..2
__device__ __forceinline__ bool ray_triangle(const RayPrep& r, const float4& q1, const float4& q2, const float4& q3,
                                             float tCurrent, float& t, float& u, float& v) {
    const Float3 A{q1.x - r.org.x, q1.y - r.org.y, q1.z - r.org.z};
    const Float3 B{q2.x - r.org.x, q2.y - r.org.y, q2.z - r.org.z};
    const Float3 C{q3.x - r.org.x, q3.y - r.org.y, q3.z - r.org.z};
    // (kx, ky, kz) = rotation of (x, y, z):  kz == 2 -> (x,y,z)   kz == 0 -> (y,z,x)   kz == 1 -> (z,x,y)
    const int p0 = r.kz == 0, p1 = r.kz == 1;
    const float Akx = fsel(p0, A.y, fsel(p1, A.z, A.x)), Aky = fsel(p0, A.z, fsel(p1, A.x, A.y)), Akz = fsel(p0, A.x, fsel(p1, A.y, A.z));
    const float Bkx = fsel(p0, B.y, fsel(p1, B.z, B.x)), Bky = fsel(p0, B.z, fsel(p1, B.x, B.y)), Bkz = fsel(p0, B.x, fsel(p1, B.y, B.z));
    const float Ckx = fsel(p0, C.y, fsel(p1, C.z, C.x)), Cky = fsel(p0, C.z, fsel(p1, C.x, C.y)), Ckz = fsel(p0, C.x, fsel(p1, C.y, C.z));
    const float Ax = Akx - r.Sx * Akz;
    const float Ay = Aky - r.Sy * Akz;
    const float Bx = Bkx - r.Sx * Bkz;
    const float By = Bky - r.Sy * Bkz;
    const float Cx = Ckx - r.Sx * Ckz;
    const float Cy = Cky - r.Sy * Ckz;
    float U = Cx * By - Cy * Bx;
    float V = Ax * Cy - Ay * Cx;
    float W = Bx * Ay - By * Ax;
    if (U == 0.0f || V == 0.0f || W == 0.0f) {
        const double CxBy = (double)Cx * (double)By, CyBx = (double)Cy * (double)Bx; U = (float)(CxBy - CyBx);
        const double AxCy = (double)Ax * (double)Cy, AyCx = (double)Ay * (double)Cx; V = (float)(AxCy - AyCx);
        const double BxAy = (double)Bx * (double)Ay, ByAx = (double)By * (double)Ax; W = (float)(BxAy - ByAx);
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
    const float det = U + V + W;
    if (det == 0.0f) return false;
    const float Az = r.Sz * Akz, Bz = r.Sz * Bkz, Cz = r.Sz * Ckz;
    const float T = U * Az + V * Bz + W * Cz;
    const uint32_t ds = __float_as_uint(det) & 0x80000000u;
    const float Ts = xor_sign(T, ds);
    if ((Ts < 0.0f) || (Ts > tCurrent * xor_sign(det, ds))) return false;
    const float rcp = 1.0f / det;
    u = U * rcp; v = V * rcp; t = T * rcp;
    return true;
}

// packed traversal link: [31] isLeaf  [30] isBlas  isBlas ? [29:10] blasOffset [9:0] idx : [29:0] idx
constexpr uint32_t kLinkLeaf = 0x80000000u, kLinkBlas = 0x40000000u;

struct TraverseResult {
    float t;            // closest accepted t (kRayMax when none)
    int tri;            // objectIdx, -1 when none
    float u, v;         // uv as TraverseBvh leaves it (updated by every watertight acceptance, ties included)
    float ua, va;       // uv at the moment the closest hit was accepted (feeds the interpolated normal)
    uint32_t nodeVisits, leafTests;
};

// One ray's traversal state.  `step` advances it by exactly one iteration of the reference loop (src/traverse.h:132-250):
// one inner-node visit, one TLAS-leaf -> BLAS-root transition, or one triangle test (+ the pop that follows a leaf or a
// double miss).  The same step function drives the plain per-thread loop (traverse_bvh) and the persistent wavefront
// kernel with lane refill (pathtrace.cu), so both produce the reference's per-ray sequence.
struct TravState {
    uint32_t cur; float curT; int top; int it;
    float t; int tri; float u, v, ua, va;
    uint32_t nodeVisits, leafTests;
};

__device__ __forceinline__ void trav_init(TravState& s) {
    s.cur = 0u; s.curT = -FLT_MAX; s.top = -1; s.it = 0;         // TLAS node 0
    s.t = kRayMax; s.tri = -1; s.u = 0.0f; s.v = 0.0f; s.ua = 0.0f; s.va = 0.0f;
    s.nodeVisits = 0; s.leafTests = 0;
}

// Advances the traversal by one reference iteration — or by two when the current link is a TLAS leaf: the
// TLAS-leaf -> BLAS-root transition (src/traverse.h:191-197) is its own iteration in the reference (and is counted
// as such in s.it against maxLoop) but costs three instructions, so the BLAS root is visited in the same physical step.
// Returns true when the traversal is finished (stack empty, any-hit satisfied, or iteration cap reached).
template <bool COUNT>
__device__ __forceinline__ bool trav_step(const SceneView& sc, const RayPrep& r, TravState& s, int anyHit, int cap, uint32_t numTri, int maxLoop,
                                          uint32_t* stackLink, float* stackT, const int stride) {
    bool pop = false;
    uint32_t cur = s.cur;
    if ((cur & (kLinkLeaf | kLinkBlas)) == kLinkLeaf) {
        cur = kLinkBlas | ((cur & 0x3FFFFFFFu) << 10);          // root of BLAS b = node 1024*b
        s.cur = cur;
        if (++s.it >= maxLoop) return true;
    }
    if (cur & kLinkLeaf) {
        const uint32_t loadIdx = cur & 0x3FFFFFFFu;          // blasOffset*1024 + idx
        float4 q1 = make_float4(0, 0, 0, 0), q2 = q1, q3 = q1;
        if (loadIdx < numTri) {                               // SAFE_LOAD (src/traverse.h:155)
            const float4* row = sc.triRows + (size_t)loadIdx * 8;
            q1 = __ldg(row); q2 = __ldg(row + 1); q3 = __ldg(row + 2);
        }
        if (COUNT) s.leafTests++;
        float tt, uu, vv;
        if (ray_triangle(r, q1, q2, q3, s.t, tt, uu, vv)) {
            s.u = uu; s.v = vv;
            if (tt < s.t) {
                s.t = tt; s.tri = (int)loadIdx; s.ua = uu; s.va = vv;
                if (anyHit) return true;
            }
        }
        pop = true;
    } else {
        float4 n0 = make_float4(0, 0, 0, 0), n1 = n0, n2 = n0, n3 = n0;
        const uint32_t idx = cur & 0x3FFFFFFFu;
        {   // one load sequence for both levels: base pointer and SAFE_LOAD limit are selected, not branched on (src/traverse.h:205-209)
            const bool blas = (cur & kLinkBlas) != 0u;
            const float4* row = (blas ? sc.nodeRows : sc.tlasRows) + (size_t)idx * 4;
            if (idx < (blas ? numTri : sc.numBatches)) { n0 = __ldg(row); n1 = __ldg(row + 1); n2 = __ldg(row + 2); n3 = __ldg(row + 3); }
        }
        if (COUNT) s.nodeVisits++;
        float t1, t2;
        const bool i1 = ray_box(r, n1.x, n1.y, n1.z, n0.x, n0.y, n0.z, t1);           // left : min = row1.xyz, max = row0.xyz
        const bool i2 = ray_box(r, n0.w, n1.w, n2.w, n2.x, n2.y, n2.z, t2);           // right: min = (row0.w,row1.w,row2.w), max = row2.xyz
        const uint32_t up = (cur & kLinkBlas) ? (cur & (kLinkBlas | 0x3FFFFC00u)) : 0u;   // keep isBlas + blasOffset
        // isLeftLeaf / isRightLeaf are written as exactly 0 or 1 by the build (bvh_block.cuh: block_write_nodes) -> bit 31 by a shift
        const uint32_t L = up | __float_as_uint(n3.x) | (__float_as_uint(n3.z) << 31);
        const uint32_t R = up | __float_as_uint(n3.y) | (__float_as_uint(n3.w) << 31);
        // near / far by selects instead of a four-way branch: the right child goes first when only it is hit, or when both are hit
        // and !(t1 < t2) (src/traverse.h:214-243: "t1 < t2 ? left first : right first")
        const bool both = i1 && i2;
        const bool rightFirst = both ? !(t1 < t2) : !i1;
        if (i1 || i2) {
            s.cur = rightFirst ? R : L; s.curT = rightFirst ? t2 : t1;
            if (both && s.top < cap - 1) {                                             // dropped when full (src/traverse.h:32-40)
                ++s.top; stackLink[s.top * stride] = rightFirst ? L : R; stackT[s.top * stride] = rightFirst ? t1 : t2;
            }
        } else pop = true;
    }
    if (pop) {                                                                         // TestForFinish (src/traverse.h:88-105)
        do {
            if (s.top < 0) return true;
            s.cur = stackLink[s.top * stride]; s.curT = stackT[s.top * stride]; --s.top;
        } while (s.curT > s.t);
    }
    return ++s.it >= maxLoop;
}

template <int STACK, bool COUNT>
__device__ __forceinline__ TraverseResult traverse_bvh(const SceneView& sc, const RayPrep& r, int anyHit,
                                                       uint32_t* stackLink, float* stackT, const int stride) {
    TravState s;
    trav_init(s);
    const int cap = sc.stackSize < STACK ? sc.stackSize : STACK;
    const uint32_t numTri = sc.numBatches * kBatchSize;
    while (!trav_step<COUNT>(sc, r, s, anyHit, cap, numTri, sc.maxLoop, stackLink, stackT, stride)) {}
    TraverseResult res;
    res.t = s.t; res.tri = s.tri; res.u = s.u; res.v = s.v; res.ua = s.ua; res.va = s.va;
    res.nodeVisits = s.nodeVisits; res.leafTests = s.leafTests;
    return res;
}

}  // namespace rtrt
