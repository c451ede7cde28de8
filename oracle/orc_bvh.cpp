// ORACLE — TEST INFRASTRUCTURE ONLY (parity pinned against oracle/_ref; see orc_types.h header).
// Stage 1: two-level LBVH build, restating
//   src/updateGeometry.cuh:13-27   MortonCode3D
//   src/updateGeometry.cuh:65-262  UpdateSceneGeometry<256,4>
//   src/radixSort.cuh:21-246       RadixSort<256,4>      (semantics: stable sort of 1024 keys)
//   src/buildBVH.cuh:8-271         LCP + BuildLBVH<256,4>
//   src/updateGeometry.cuh:267-364 UpdateTLAS<256,4,1024>
//   src/bvh.cu:7-97                BuildBvhLevel1/2 sequencing
//   src/init.cu:103-130            padding / triCountArray
#include "orc_types.h"
#include <vector>
#include <thread>
#include <algorithm>
#include <numeric>

namespace orc {

// src/updateGeometry.cuh:13-27
u32 MortonCode3D(u32 x, u32 y, u32 z) {
    x = (x | (x << 16)) & 0x030000FF; x = (x | (x << 8)) & 0x0300F00F; x = (x | (x << 4)) & 0x030C30C3; x = (x | (x << 2)) & 0x09249249;
    y = (y | (y << 16)) & 0x030000FF; y = (y | (y << 8)) & 0x0300F00F; y = (y | (y << 4)) & 0x030C30C3; y = (y | (y << 2)) & 0x09249249;
    z = (z | (z << 16)) & 0x030000FF; z = (z | (z << 8)) & 0x0300F00F; z = (z | (z << 4)) & 0x030C30C3; z = (z | (z << 2)) & 0x09249249;
    return x | (y << 1) | (z << 2);
}

// src/updateGeometry.cuh:254-258 : (c - min) / (max - min), *1023, CUDA float->uint
static u32 morton_of(F3 c, const AABB& box) {
    F3 u = (c - box.min) / (box.max - box.min);
    return MortonCode3D(cuda_f2u(u.x * 1023.0f), cuda_f2u(u.y * 1023.0f), cuda_f2u(u.z * 1023.0f));
}

static inline int clz32(u32 v) { return v == 0 ? 32 : __builtin_clz(v); }

// src/buildBVH.cuh:8-14 LCP.  Deviation (documented in DESIGN.md): equal keys are tie-broken by
// position (Karras 2012 §4), which yields the reference's tree whenever no key occurs 3+ times
// inside one segment and keeps the hierarchy a valid tree when it does (the reference's is not).
static inline int delta(const u32* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return 0;
    u32 x = keys[i] ^ keys[j];
    if (x != 0) return clz32(x);
    return 32 + clz32((u32)i ^ (u32)j);
}

// Karras hierarchy for one segment: src/buildBVH.cuh:60-134
// keys: n sorted keys; reorder: leaf payload per sorted position; nodes: n-1 nodes out.
static void karras_emit(const u32* keys, const u32* reorder, int n, BVHNode* nodes) {
    for (int i = 0; i < n - 1; ++i) {
        int dl = delta(keys, n, i, i - 1), dr = delta(keys, n, i, i + 1);
        int d = (dr - dl) >= 0 ? 1 : -1;
        int dmin = delta(keys, n, i, i - d);
        int lmax = 2;
        while (delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
        int l = 0;
        for (int t = lmax / 2; t >= 1; t /= 2)
            if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
        int j = i + l * d;
        int dnode = delta(keys, n, i, j);
        int s = 0;
        // t = ceil(l/2), ceil(l/4), ..., 1 (the reference keeps re-testing t=1 until `div`
        // overflows; those extra probes never succeed on sorted keys)
        for (int div = 2;; div *= 2) {
            int t = (l + div - 1) / div;
            if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
            if (t <= 1) break;
        }
        int gamma = i + s * d + std::min(d, 0);
        BVHNode& nd = nodes[i];
        if (std::min(i, j) == gamma) { nd.isLeftLeaf = 1; nd.idxLeft = reorder[gamma]; }
        else                         { nd.isLeftLeaf = 0; nd.idxLeft = (u32)gamma; }
        if (std::max(i, j) == gamma + 1) { nd.isRightLeaf = 1; nd.idxRight = reorder[gamma + 1]; }
        else                             { nd.isRightLeaf = 0; nd.idxRight = (u32)(gamma + 1); }
    }
}

// bottom-up refit: src/buildBVH.cuh:138-267 (result = exact unions, order independent)
static AABB refit(BVHNode* nodes, const AABB* leafBoxes, u32 i) {
    BVHNode& nd = nodes[i];
    AABB l = nd.isLeftLeaf  ? leafBoxes[nd.idxLeft]  : refit(nodes, leafBoxes, nd.idxLeft);
    AABB r = nd.isRightLeaf ? leafBoxes[nd.idxRight] : refit(nodes, leafBoxes, nd.idxRight);
    node_set_left(nd, l);
    node_set_right(nd, r);
    return node_merged(nd);
}

// n == 1 special case: src/buildBVH.cuh:31-38 (right box is the point AABB(0,0))
static void single_leaf_node(BVHNode& nd, const AABB& box) {
    node_set_left(nd, box);
    node_set_right(nd, AABB{f3(0.0f), f3(0.0f)});
    nd.idxLeft = 0; nd.idxRight = 0; nd.isLeftLeaf = 1; nd.isRightLeaf = 1;
}

// stable sort of a segment of keys -> sorted keys (in place) + reorder  (src/radixSort.cuh)
static void segment_sort(u32* keys, u32* reorder, int n) {
    std::vector<u32> idx(n);
    std::iota(idx.begin(), idx.end(), 0u);
    std::stable_sort(idx.begin(), idx.end(), [&](u32 a, u32 b) { return keys[a] < keys[b]; });
    std::vector<u32> k(n);
    for (int i = 0; i < n; ++i) k[i] = keys[idx[i]];
    for (int i = 0; i < n; ++i) { keys[i] = k[i]; reorder[i] = idx[i]; }
}

struct BuildArgs {
    const float* pos; const float* nrm; const u32* idx; u32 V, N;
    Triangle* triangles; AABB* aabbs; u32* mortonPre; u32* morton; u32* reorderIdx; BVHNode* nodes;
};

static inline F3 ld3(const float* p, u32 i) { return F3{p[3 * i], p[3 * i + 1], p[3 * i + 2]}; }

// one batch: UpdateSceneGeometry + RadixSort + BuildLBVH for block b
static void build_batch(const BuildArgs& a, u32 b, u32 count) {
    const u32 base = b * kBatch;
    const u32 nPadded4 = (a.N + 3u) & ~3u;                 // src/init.cu:103-115 (indices padded with 0)
    u32 active = (count + 3u) & ~3u;                       // threads with 4t <= count-1 handle 4 items each (:93)
    AABB batchBox = aabb_empty();
    std::vector<F3> centre(active);
    for (u32 i = 0; i < active; ++i) {
        u32 g = base + i;
        u32 i0 = 0, i1 = 0, i2 = 0;
        if (g < a.N) { i0 = a.idx[3 * g]; i1 = a.idx[3 * g + 1]; i2 = a.idx[3 * g + 2]; }
        (void)nPadded4;
        Triangle t; std::memset(&t, 0, sizeof(t));
        t.v1 = ld3(a.pos, i0); t.v2 = ld3(a.pos, i1); t.v3 = ld3(a.pos, i2);          // :131-133
        t.n1 = ld3(a.nrm, i0); t.n2 = ld3(a.nrm, i1); t.n3 = ld3(a.nrm, i2);          // :135-137
        a.triangles[g] = t;
        F3 mn = min3f(t.v1, min3f(t.v2, t.v3));                                       // :172
        F3 mx = max3f(t.v1, max3f(t.v2, t.v3));                                       // :173
        F3 diff = max3f(mx - mn, MachineEpsilon() * mx);                              // :176
        mx = mn + diff;                                                               // :177
        a.aabbs[g] = AABB{mn, mx};
        centre[i] = (t.v1 + t.v2 + t.v3) / 3.0f;                                      // :180
        batchBox.min = min3f(batchBox.min, mn);
        batchBox.max = max3f(batchBox.max, mx);
    }
    for (u32 i = 0; i < kBatch; ++i)
        a.morton[base + i] = (i < active) ? morton_of(centre[i], batchBox) : 0xFFFFFFFFu;  // src/kernel.cu:279
    if (a.mortonPre) std::memcpy(a.mortonPre + base, a.morton + base, kBatch * sizeof(u32));
    segment_sort(a.morton + base, a.reorderIdx + base, kBatch);
    BVHNode* nodes = a.nodes + base;
    if (count == 1) { single_leaf_node(nodes[0], a.aabbs[base]); return; }
    karras_emit(a.morton + base, a.reorderIdx + base, (int)count, nodes);
    refit(nodes, a.aabbs + base, 0);
}

}  // namespace orc

using namespace orc;

extern "C" {

// Sizes: B = ceil(N/1024); triangles/aabbs/morton*/reorderIdx/nodes hold B*1024 entries (zero/FF filled
// where the reference leaves memory untouched); tlas* hold Bp = ceil(B/1024)*1024 entries.
// mortonPre / tlasMortonPre may be NULL.
int orc_build_bvh(const float* pos, const float* nrm, const u32* idx, u32 V, u32 N,
                  void* triangles, void* aabbs, u32* mortonPre, u32* morton, u32* reorderIdx, void* nodes,
                  void* tlasAabbs, u32* tlasMortonPre, u32* tlasMorton, u32* tlasReorderIdx, void* tlasNodes,
                  int threads) {
    if (N == 0) return -1;
    const u32 B = (N + kBatch - 1) / kBatch;
    BuildArgs a{pos, nrm, idx, V, N, (Triangle*)triangles, (AABB*)aabbs, mortonPre, morton, reorderIdx, (BVHNode*)nodes};
    std::memset(triangles, 0, (size_t)B * kBatch * sizeof(Triangle));
    std::memset(aabbs, 0, (size_t)B * kBatch * sizeof(AABB));
    std::memset(nodes, 0, (size_t)B * kBatch * sizeof(BVHNode));
    if (threads < 1) threads = 1;
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
        pool.emplace_back([&, t]() {
            for (u32 b = t; b < B; b += threads) {
                u32 count = (b == B - 1) ? (N - kBatch * (B - 1)) : kBatch;     // src/init.cu:129-130
                build_batch(a, b, count);
            }
        });
    for (auto& th : pool) th.join();

    // ---- TLAS: src/updateGeometry.cuh:267-364, then sort + Karras over B items ----
    const u32 Bp = (B + kBatch - 1) / kBatch * kBatch;
    AABB* tb = (AABB*)tlasAabbs; BVHNode* tn = (BVHNode*)tlasNodes; const BVHNode* bn = (const BVHNode*)nodes;
    std::memset(tb, 0, (size_t)Bp * sizeof(AABB));
    std::memset(tn, 0, (size_t)Bp * sizeof(BVHNode));
    AABB scene = aabb_empty();
    std::vector<F3> centre(B);
    for (u32 k = 0; k < B; ++k) {
        AABB m = node_merged(bn[(size_t)kBatch * k]);                            // :294-296
        centre[k] = (m.max + m.min) / 2.0f;                                      // :297
        tb[k] = m;
        // UpdateTLAS reduces its 256 per-thread boxes with strides 128 and 64 and then lets warp 0 reduce entries 0..31 only
        // (src/updateGeometry.cuh:313-330; the "tid < 32: merge tid + 32" step of UpdateSceneGeometry :219-223 is missing), so
        // the boxes of threads 32..63, 96..127, 160..191 and 224..255 never reach the scene box: batches with k mod 256 >= 128
        // are left out.  Reproduced for every B the reference can run (< 1024); the multi-CTA generalisation uses all batches.
        if (B >= kBatch || (k & 255u) < 128u) { scene.min = min3f(scene.min, m.min); scene.max = max3f(scene.max, m.max); }
    }
    for (u32 k = 0; k < Bp; ++k) tlasMorton[k] = (k < B) ? morton_of(centre[k], scene) : 0xFFFFFFFFu;
    if (tlasMortonPre) std::memcpy(tlasMortonPre, tlasMorton, Bp * sizeof(u32));
    segment_sort(tlasMorton, tlasReorderIdx, (int)Bp);
    if (B == 1) single_leaf_node(tn[0], tb[0]);
    else { karras_emit(tlasMorton, tlasReorderIdx, (int)B, tn); refit(tn, tb, 0); }
    return 0;
}

}  // extern "C"
