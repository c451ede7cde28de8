// ORACLE — TEST INFRASTRUCTURE ONLY (parity pinned against oracle/_ref; see orc_types.h header).
// Stage 2 core: two-level stack traversal + watertight triangle + robust box test, restating
//   src/geometry.cuh:375-472   max_dim / sign_mask / xorf / RayTriangleWatertight
//   src/geometry.cuh:497-616   RayBoxIntersectionHelper / RayAABBIntersect / RayAabbPairIntersect
//   src/geometry.cuh:252-262   GetRayPlaneIntersectPoint
//   src/traverse.h:9-253       BvhNodeStack / TestForFinish / TraverseBvh
#pragma once
#include "orc_types.h"

namespace orc {

static inline int max_dim(F3 d) {                         // src/geometry.cuh:375-387
    if (d.x > d.y) { if (d.z > d.x) return 2; else return 0; }
    else           { if (d.z > d.y) return 2; else return 1; }
}
static inline int sign_mask(float a) { int b; std::memcpy(&b, &a, 4); return b & (int)0x80000000; }
static inline float xorf(float a, int b) { int c; std::memcpy(&c, &a, 4); c ^= b; float d; std::memcpy(&d, &c, 4); return d; }

// src/geometry.cuh:406-472
static inline bool RayTriangleWatertight(const Ray& ray, const Triangle& tri, float tCurrentHit,
                                         float& t, float& u, float& v, float& e) {
    F3 dir = ray.dir, org = ray.orig;
    int kz = max_dim(abs3(dir));
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    if (dir[kz] < 0.0f) { int tmp = kx; kx = ky; ky = tmp; }
    float Sx = dir[kx] / dir[kz];
    float Sy = dir[ky] / dir[kz];
    float Sz = 1.0f / dir[kz];
    const F3 A = tri.v1 - org, B = tri.v2 - org, C = tri.v3 - org;
    const float Ax = A[kx] - Sx * A[kz];
    const float Ay = A[ky] - Sy * A[kz];
    const float Bx = B[kx] - Sx * B[kz];
    const float By = B[ky] - Sy * B[kz];
    const float Cx = C[kx] - Sx * C[kz];
    const float Cy = C[ky] - Sy * C[kz];
    float U = Cx * By - Cy * Bx;
    float V = Ax * Cy - Ay * Cx;
    float W = Bx * Ay - By * Ax;
    if (U == 0.0f || V == 0.0f || W == 0.0f) {
        double CxBy = (double)Cx * (double)By, CyBx = (double)Cy * (double)Bx; U = (float)(CxBy - CyBx);
        double AxCy = (double)Ax * (double)Cy, AyCx = (double)Ay * (double)Cx; V = (float)(AxCy - AyCx);
        double BxAy = (double)Bx * (double)Ay, ByAx = (double)By * (double)Ax; W = (float)(BxAy - ByAx);
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
    float det = U + V + W;
    if (det == 0.0f) return false;
    const float Az = Sz * A[kz], Bz = Sz * B[kz], Cz = Sz * C[kz];
    const float T = U * Az + V * Bz + W * Cz;
    int det_sign = sign_mask(det);
    if ((xorf(T, det_sign) < 0.0f) || (xorf(T, det_sign) > tCurrentHit * xorf(det, det_sign))) return false;
    const float rcpDet = 1.0f / det;
    u = U * rcpDet; v = V * rcpDet; t = T * rcpDet;
    e = ErrGamma(16) * fabsf(t);
    return true;
}

struct BoxHelper {                                          // src/geometry.cuh:497-517
    int nearX, nearY, nearZ, farX, farY, farZ;
    float org_near_x, org_near_y, org_near_z, org_far_x, org_far_y, org_far_z;
    float rdir_near_x, rdir_near_y, rdir_near_z, rdir_far_x, rdir_far_y, rdir_far_z;
};

// src/geometry.cuh:519-583
static inline BoxHelper CreateRayBoxIntersectionHelper(const Ray& ray, const AABB& sceneBox, F3 inv) {
    int kz = max_dim(abs3(ray.dir));
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    if (ray.dir[kz] < 0.0f) { int tmp = kx; kx = ky; ky = tmp; }
    int nearID[3] = {0, 1, 2}, farID[3] = {3, 4, 5};
    int nearX = nearID[kx], farX = farID[kx];
    int nearY = nearID[ky], farY = farID[ky];
    int nearZ = nearID[kz], farZ = farID[kz];
    if (ray.dir[kx] < 0.0f) { int tmp = nearX; nearX = farX; farX = tmp; }
    if (ray.dir[ky] < 0.0f) { int tmp = nearY; nearY = farY; farY = tmp; }
    if (ray.dir[kz] < 0.0f) { int tmp = nearZ; nearZ = farZ; farZ = tmp; }
    const float eps = 5.0f * TwoPowerMinus24();
    F3 lower = Dn(abs3(ray.orig - sceneBox.min));
    F3 upper = Up(abs3(ray.orig - sceneBox.max));
    float max_z = rmax(lower[kz], upper[kz]);
    float err_near_x = Up(lower[kx] + max_z);
    float err_near_y = Up(lower[ky] + max_z);
    float org_near_x = up(ray.orig[kx] + Up(eps * err_near_x));
    float org_near_y = up(ray.orig[ky] + Up(eps * err_near_y));
    float org_near_z = ray.orig[kz];
    float err_far_x = Up(upper[kx] + max_z);
    float err_far_y = Up(upper[ky] + max_z);
    float org_far_x = dn(ray.orig[kx] - Up(eps * err_far_x));
    float org_far_y = dn(ray.orig[ky] - Up(eps * err_far_y));
    float org_far_z = ray.orig[kz];
    if (ray.dir[kx] < 0.0f) { float tmp = org_near_x; org_near_x = org_far_x; org_far_x = tmp; }
    if (ray.dir[ky] < 0.0f) { float tmp = org_near_y; org_near_y = org_far_y; org_far_y = tmp; }
    BoxHelper h;
    h.nearX = nearX; h.nearY = nearY; h.nearZ = nearZ; h.farX = farX; h.farY = farY; h.farZ = farZ;
    h.org_near_x = org_near_x; h.org_near_y = org_near_y; h.org_near_z = org_near_z;
    h.org_far_x = org_far_x; h.org_far_y = org_far_y; h.org_far_z = org_far_z;
    h.rdir_near_x = Dn(Dn(inv[kx])); h.rdir_near_y = Dn(Dn(inv[ky])); h.rdir_near_z = Dn(Dn(inv[kz]));
    h.rdir_far_x = Up(Up(inv[kx])); h.rdir_far_y = Up(Up(inv[ky])); h.rdir_far_z = Up(Up(inv[kz]));
    return h;
}

// src/geometry.cuh:585-603
static inline bool RayAABBIntersect(const AABB& b, const BoxHelper& h, float& tNear, float& tFar) {
    float tNearX = (b[h.nearX] - h.org_near_x) * h.rdir_near_x;
    float tNearY = (b[h.nearY] - h.org_near_y) * h.rdir_near_y;
    float tNearZ = (b[h.nearZ] - h.org_near_z) * h.rdir_near_z;
    float tFarX = (b[h.farX] - h.org_far_x) * h.rdir_far_x;
    float tFarY = (b[h.farY] - h.org_far_y) * h.rdir_far_y;
    float tFarZ = (b[h.farZ] - h.org_far_z) * h.rdir_far_z;
    tNear = rmax3(tNearX, tNearY, tNearZ);
    tFar = rmin3(tFarX, tFarY, tFarZ);
    bool hit = tNear <= tFar && (tFar > 0);
    tNear = rmax(tNear, 0.0f);
    return hit;
}

// src/geometry.cuh:252-262
static inline F3 GetRayPlaneIntersectPoint(F3 n, float w, const Ray& ray, float t, float& pError) {
    F3 p = ray.orig + ray.dir * t;
    F3 pc = p - (dot(n, p) + w) * n;
    pError = getmax(abs3(pc)) * ErrGamma(6);
    return pc;
}

struct Scene {
    const Triangle* triangles; const BVHNode* nodes; const BVHNode* tlas;
    u32 numBatches;          // B
    int stackSize;           // 16 = reference (src/traverse.h:28)
    int maxLoop;             // 1024 = reference (src/traverse.h:132)
};

struct HitRecord {
    float t; F2 uv; int objectIdx; F3 normal; F3 fakeNormal; F3 point; float rayOffset; float errorP, errorT;
    u32 nodeVisits, leafTests;   // counters for the secondary metric
};

struct StackEntry { u32 idx, blasOffset; int isBlas, isLeaf; float t; };

// src/traverse.h:107-253.  anyHit != 0: stop at the first accepted triangle.
static inline void TraverseBvh(const Scene& sc, const Ray& ray, F3 inv, HitRecord& r, int anyHit) {
    const BVHNode top = sc.tlas[0];
    AABB sceneBox = node_merged(top);
    BoxHelper h = CreateRayBoxIntersectionHelper(ray, sceneBox, inv);
    StackEntry stack[64];
    int top_ = -1;
    const int cap = sc.stackSize > 64 ? 64 : sc.stackSize;
    StackEntry cur{0, 0, 0, 0, -FLT_MAX};
    const int trianglesSize = (int)(sc.numBatches * kBatch);
    auto finish = [&]() -> bool {                            // TestForFinish src/traverse.h:88-105
        do {
            if (top_ < 0) return true;
            cur = stack[top_--];
        } while (cur.t > r.t);
        return false;
    };
    for (int it = 0; it < sc.maxLoop; ++it) {
        if (cur.isLeaf) {
            if (cur.isBlas) {
                int loadIdx = (int)(cur.blasOffset * kBatch + cur.idx);
                Triangle tri; std::memset(&tri, 0, sizeof(tri));
                if (loadIdx >= 0 && loadIdx < trianglesSize) tri = sc.triangles[loadIdx];   // SAFE_LOAD
                r.leafTests++;
                float tt;
                bool hit = RayTriangleWatertight(ray, tri, r.t, tt, r.uv.x, r.uv.y, r.errorT);
                float t_temp = hit ? tt : RayMax;                                          // geometry.cuh:474-495
                if (t_temp < r.t) {
                    r.t = t_temp;
                    r.objectIdx = loadIdx;
                    r.normal = normalize(cross(tri.v2 - tri.v1, tri.v3 - tri.v1));
                    r.point = GetRayPlaneIntersectPoint(r.normal, -dot(r.normal, tri.v1), ray, r.t, r.errorP);
                    r.rayOffset = r.errorT + r.errorP;
                    F3 n1 = normalize(tri.n1), n2 = normalize(tri.n2), n3 = normalize(tri.n3);
                    r.fakeNormal = normalize(n3 * (1.0f - r.uv.x - r.uv.y) + n1 * r.uv.x + n2 * r.uv.y);
                    if (anyHit) break;
                }
                if (finish()) break;
            } else {
                cur.isLeaf = 0; cur.isBlas = 1; cur.blasOffset = cur.idx; cur.idx = 0;
            }
        } else {
            BVHNode nd; std::memset(&nd, 0, sizeof(nd));
            if (cur.isBlas) {
                long li = (long)cur.blasOffset * kBatch + cur.idx;
                if (li >= 0 && li < (long)trianglesSize) nd = sc.nodes[li];
            } else {
                if (cur.idx < sc.numBatches) nd = sc.tlas[cur.idx];
            }
            r.nodeVisits++;
            float t1, t2, f1, f2;
            bool i1 = RayAABBIntersect(node_left(nd), h, t1, f1);
            bool i2 = RayAABBIntersect(node_right(nd), h, t2, f2);
            if (!i1 && !i2) { if (finish()) break; }
            else if (i1 && !i2) { cur.idx = nd.idxLeft; cur.isLeaf = (int)nd.isLeftLeaf; cur.t = t1; }
            else if (!i1 && i2) { cur.idx = nd.idxRight; cur.isLeaf = (int)nd.isRightLeaf; cur.t = t2; }
            else {
                StackEntry far;
                if (t1 < t2) {
                    far = StackEntry{nd.idxRight, cur.blasOffset, cur.isBlas, (int)nd.isRightLeaf, t2};
                    cur.idx = nd.idxLeft; cur.isLeaf = (int)nd.isLeftLeaf; cur.t = t1;
                } else {
                    far = StackEntry{nd.idxLeft, cur.blasOffset, cur.isBlas, (int)nd.isLeftLeaf, t1};
                    cur.idx = nd.idxRight; cur.isLeaf = (int)nd.isRightLeaf; cur.t = t2;
                }
                if (top_ < cap - 1) stack[++top_] = far;   // push on a full stack is dropped (traverse.h:32-40)
            }
        }
    }
}

// src/traverse.cuh:86-121 : state initialisation done by RaySceneIntersect before TraverseBvh
static inline HitRecord init_hit() {
    HitRecord r;
    r.t = RayMax; r.uv = F2{0, 0}; r.objectIdx = -1; r.normal = f3(0.0f); r.fakeNormal = f3(0.0f);
    r.point = f3(RayMax); r.rayOffset = 1e-7f; r.errorP = 1e-7f; r.errorT = 1e-7f;
    r.nodeVisits = 0; r.leafTests = 0;
    return r;
}

}  // namespace orc
