// Builds the wide traversal layout (wide_bvh.cuh) from the reference-layout arrays of a finished two-level LBVH build:
//   k_wide_blas  one CTA per 1024-triangle batch: collapses the batch's binary radix tree (BVHNode[], src/bvhNode.cuh:5-13) into
//                8-wide nodes, top-down and level-synchronously; 8 lanes own the 8 child slots of one wide node;
//   k_wide_tlas  one CTA: the same collapse over the top-level tree; a top-level leaf (= batch) becomes an inner child whose node
//                is a copy of that batch's wide root, so the two levels form ONE wide tree.
// The binary tree is never re-derived: ranges of sorted positions are carried down with Karras' numbering (the inner left / right
// children of the node that splits at gamma are nodes gamma / gamma + 1, bvh_block.cuh), which gives every subtree's triangle count
// for free.  Collapse rule (Ylitie et al. 2017, section 4.1, greedy form): start from the two children of the binary node, keep
// opening the inner child with the largest surface area until there are 8 children; subtrees of <= kWideMaxLeaf triangles (1 since round 2, the format allows 3) are leaves.
// Nodes are allocated breadth-first from a per-batch region (inner children of a node are consecutive), triangles are copied
// (positions only, original index in row0.w) in the order the leaves were allocated, so a node's leaf triangles are consecutive too.
#include "wide_bvh.cuh"
#include <cuda_fp16.h>
#include "context.h"

namespace rtrt {

#ifndef RTRT_WB_THREADS
#define RTRT_WB_THREADS 256
#endif
constexpr int kWbThreads = RTRT_WB_THREADS;              // k_wide_blas: groups of 8 lanes, one wide node per group and round
constexpr int kWtThreads = 1024;                         // k_wide_tlas: 128 groups
#ifndef RTRT_WIDE_DISSOLVE
#define RTRT_WIDE_DISSOLVE 1
#endif
#ifndef RTRT_WIDE_PACK_LEAVES
#define RTRT_WIDE_PACK_LEAVES 0      // measured, see wide_process: fewer nodes (321 k -> 248 k) but MORE visits per ray (24.1 -> 26.9), bounce round 0.85 -> 0.91 ms
#endif
constexpr uint32_t kCopyFlag = 0x80000000u;              // queue entry: "copy the wide root of batch (link & ~flag)"
enum { K_EMPTY = 0, K_INNER = 1, K_LEAF = 2, K_BATCH = 3 };

struct WideEntry {                                        // one pending wide node: the binary subtree it is made from
    uint32_t link, f, l;                                  // binary node index (or batch | kCopyFlag), range of sorted positions
    float lo[3], hi[3];                                   // its box (exact union of the leaves: min / max are exact)
    uint32_t pad;
};
static_assert(sizeof(WideEntry) == 40, "WideEntry");

struct WideBuildParams {
    const float4* __restrict__ binNodes;                  // BLAS: BVHNode[B*1024] as 4 float4 rows; TLAS: the top-level nodes
    const uint32_t* __restrict__ reorderIdx;              // BLAS: sorted position -> batch-local triangle id
    const float4* __restrict__ triRows;                   // Triangle[] as 8 float4 rows
    float4* nodes; float4* tris;                          // outputs
    uint32_t numTris, numBatches;
    uint32_t tlasBase;                                    // first node index of the top-level region (= numBatches * kWideRegion)
    uint32_t tlasCap;                                     // node slots of the top-level region
    WideEntry* tlasQueue;                                 // global-memory queue of k_wide_tlas (tlasCap entries)
    uint32_t* stat;                                       // [0] error flags  [1] nodes written  [2] top-level nodes
};

struct Slot { int kind; uint32_t idx, f, l; float lo[3], hi[3]; };

__device__ __forceinline__ float box_area(const Slot& s) {
    const float dx = s.hi[0] - s.lo[0], dy = s.hi[1] - s.lo[1], dz = s.hi[2] - s.lo[2];
    return dx * dy + dy * dz + dz * dx;
}

// the two children of binary node `rec` (4 rows) whose range is [f, l]
template <bool TLAS>
__device__ __forceinline__ void child_of(const float4 (&rec)[4], uint32_t f, uint32_t l, bool right, Slot& s) {
    const uint32_t idxL = __float_as_uint(rec[3].x), idxR = __float_as_uint(rec[3].y);
    const bool leafL = __float_as_uint(rec[3].z) != 0u, leafR = __float_as_uint(rec[3].w) != 0u;
    const uint32_t gamma = leafL ? f : idxL;
    if (!right) {
        s.idx = idxL; s.f = f; s.l = gamma;
        s.lo[0] = rec[1].x; s.lo[1] = rec[1].y; s.lo[2] = rec[1].z; s.hi[0] = rec[0].x; s.hi[1] = rec[0].y; s.hi[2] = rec[0].z;
        s.kind = TLAS ? (leafL ? K_BATCH : K_INNER) : ((s.l - s.f + 1u <= (uint32_t)kWideMaxLeaf) ? K_LEAF : K_INNER);
    } else {
        s.idx = idxR; s.f = gamma + 1u; s.l = l;
        s.lo[0] = rec[0].w; s.lo[1] = rec[1].w; s.lo[2] = rec[2].w; s.hi[0] = rec[2].x; s.hi[1] = rec[2].y; s.hi[2] = rec[2].z;
        s.kind = TLAS ? (leafR ? K_BATCH : K_INNER) : ((s.l - s.f + 1u <= (uint32_t)kWideMaxLeaf) ? K_LEAF : K_INNER);
    }
}

__device__ __forceinline__ void load_rec(const float4* __restrict__ binNodes, uint32_t idx, float4 (&rec)[4]) {
    const float4* r = binNodes + (size_t)idx * 4;
    rec[0] = __ldcg(r); rec[1] = __ldcg(r + 1); rec[2] = __ldcg(r + 2); rec[3] = __ldcg(r + 3);     // written by the build just before: L2
}

__device__ __forceinline__ uint32_t exp_byte(float extent) {          // smallest e with 2^(e-127) >= extent
    const uint32_t b = __float_as_uint(extent);
    uint32_t e = ((b >> 23) & 0xFFu) + ((b & 0x007FFFFFu) ? 1u : 0u);
    return e < 1u ? 1u : (e > 253u ? 253u : e);
}

// One wide node per 8-lane group.  Every lane of the warp runs this (valid = the group has an entry), shuffles use the full mask.
//   binNodes  binary nodes of this tree (already offset to the batch for the BLAS)
//   queue     entries of this tree's region; nodeTail / triTail its allocation counters
//   outIndex  region-local index of the node to write; regionBase / triRegion the global offsets of the region
template <bool TLAS>
__device__ __forceinline__ void wide_process(const WideBuildParams& p, const float4* __restrict__ binNodes, WideEntry* queue, uint32_t w, bool valid,
                                             uint32_t* nodeTail, uint32_t* triTail, uint16_t* triSrc, uint32_t regionBase, uint32_t regionCap,
                                             uint32_t triRegion, bool singleLeaf) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, g = lane & 7, gb = lane & ~7;
    WideEntry e{};
    if (valid) e = queue[w];
    float4* outNode = p.nodes + (size_t)(regionBase + w) * kWideStride;
    if (TLAS && valid && (e.link & kCopyFlag)) valid = false;      // top-level leaf: the node is a copy of the batch's wide root (k_wide_copy, once both trees exist)
    Slot s; s.kind = K_EMPTY; s.idx = 0; s.f = 0; s.l = 0;
    for (int a = 0; a < 3; ++a) { s.lo[a] = 0.0f; s.hi[a] = 0.0f; }
    int nSlots = 0;
    if (valid) {
        float4 rec[4];
        load_rec(binNodes, e.link, rec);
        if (g == 0) child_of<TLAS>(rec, e.f, e.l, false, s);
        if (g == 1 && !singleLeaf) child_of<TLAS>(rec, e.f, e.l, true, s);
        nSlots = singleLeaf ? 1 : 2;
    }
    // ---- open the inner child with the largest surface area until 8 slots are in use ----
    // Leaf packing (RTRT_WIDE_PACK_LEAVES, off): a node over more than 8 items (triangles / batches) only opens children that themselves hold more than 8,
    // so every maximal subtree of <= 8 items becomes ONE node whose slots are its items.  Pure largest-area opening gives 4.3 children per node
    // (3.3 triangles + 1.0 inner) on the bench scene, packing 5.3 — but packed triangles always sit one level lower, in nodes of their own, whereas the
    // mixed nodes of the greedy collapse let a ray test triangle boxes and inner boxes in one visit: measured slower (tools/gpu_r2w.sh).
#if RTRT_WIDE_PACK_LEAVES
    const bool smallNode = (e.l - e.f + 1u) <= 8u;
#endif
#pragma unroll 1
    for (int it = 0; it < 6; ++it) {
#if RTRT_WIDE_PACK_LEAVES
        float best = (valid && s.kind == K_INNER && (smallNode || (s.l - s.f + 1u) > 8u)) ? box_area(s) : -1.0f;
#else
        float best = (valid && s.kind == K_INNER) ? box_area(s) : -1.0f;
#endif
#if RTRT_WIDE_DISSOLVE
        // a child whose whole subtree fits into the slots that are still free (k items take k - 1 more slots) is opened before any larger one:
        // dissolving it costs nothing per visit (all 8 slots are tested anyway) and saves a node — and a visit — below
        if (!TLAS && best >= 0.0f && (s.l - s.f + 1u) <= 9u - (uint32_t)nSlots) best += 1.0e30f;
#endif
        int who = g;
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            const float ob = __shfl_xor_sync(full, best, o); const int ow = __shfl_xor_sync(full, who, o);
            if (ob > best || (ob == best && ow < who)) { best = ob; who = ow; }
        }
        if (!__any_sync(full, valid && best >= 0.0f && nSlots < 8)) break;     // no group of this warp has an inner child left to open
        const uint32_t widx = __shfl_sync(full, s.idx, gb + who), wf = __shfl_sync(full, s.f, gb + who), wl = __shfl_sync(full, s.l, gb + who);
        if (valid && best >= 0.0f && nSlots < 8) {
            float4 rec[4];
            load_rec(binNodes, widx, rec);
            if (g == who) child_of<TLAS>(rec, wf, wl, false, s);
            if (g == nSlots) child_of<TLAS>(rec, wf, wl, true, s);
            ++nSlots;
        }
    }
    // ---- slot assignment: child with displacement d from the node centre prefers slot s maximising sum_axis (+-)d, sign + where the
    // slot's axis bit is set (x = bit 2, y = bit 1, z = bit 0): rays visit slots in descending (slot ^ (7 - octant)) ----
    const float cx = (s.lo[0] + s.hi[0]) - (e.lo[0] + e.hi[0]), cy = (s.lo[1] + s.hi[1]) - (e.lo[1] + e.hi[1]), cz = (s.lo[2] + s.hi[2]) - (e.lo[2] + e.hi[2]);
    int mySlot = -1;
    uint32_t freeMask = 0xFFu;
    // the child's score for each of the 8 slots does not change between the rounds of the auction (only the free mask does): computed once
    float sc[8];
#pragma unroll
    for (int t = 0; t < 8; ++t) {
        float c = ((t & 4) ? cx : -cx) + ((t & 2) ? cy : -cy) + ((t & 1) ? cz : -cz);
        if (!(c == c)) c = 0.0f;                                          // a NaN box must not lose its child
        sc[t] = c;
    }
    const bool bidder = valid && s.kind != K_EMPTY;
    const int rounds = __shfl_sync(full, nSlots, gb);                     // one child is placed per round
#pragma unroll 1
    for (int it = 0; it < 8; ++it) {
        if (__all_sync(full, it >= rounds)) break;
        float best = -3.0e38f; int bs = 0;
        if (bidder && mySlot < 0) {
#pragma unroll
            for (int t = 0; t < 8; ++t)
                if (((freeMask >> t) & 1u) && sc[t] > best) { best = sc[t]; bs = t; }
        }
        int who = g; float wb = best;
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            const float ob = __shfl_xor_sync(full, wb, o); const int ow = __shfl_xor_sync(full, who, o);
            if (ob > wb || (ob == wb && ow < who)) { wb = ob; who = ow; }
        }
        const int ws = __shfl_sync(full, bs, gb + who);
        if (wb > -3.0e38f) {
            if (g == who) mySlot = ws;
            freeMask &= ~(1u << ws);
        }
    }
    // ---- move every child to the lane of its slot ----
    int src = -1;
#pragma unroll
    for (int j = 0; j < 8; ++j) { const int sj = __shfl_sync(full, mySlot, gb + j); if (sj == g) src = j; }
    const int from = gb + (src < 0 ? g : src);
    Slot c;
    c.kind = __shfl_sync(full, s.kind, from); c.idx = __shfl_sync(full, s.idx, from); c.f = __shfl_sync(full, s.f, from); c.l = __shfl_sync(full, s.l, from);
#pragma unroll
    for (int a = 0; a < 3; ++a) { c.lo[a] = __shfl_sync(full, s.lo[a], from); c.hi[a] = __shfl_sync(full, s.hi[a], from); }
    if (src < 0 || !valid) c.kind = K_EMPTY;
    // ---- allocation: inner children consecutive in slot order, leaf triangles consecutive in slot order ----
    const bool inner = c.kind == K_INNER || c.kind == K_BATCH;
    const uint32_t imask = (__ballot_sync(full, inner) >> gb) & 0xFFu;
    const uint32_t cnt = (c.kind == K_LEAF) ? (c.l - c.f + 1u) : 0u;
    uint32_t incl = cnt;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) { const uint32_t v = __shfl_up_sync(full, incl, o, 8); if (g >= o) incl += v; }
    const uint32_t triOff = incl - cnt, triTotal = __shfl_sync(full, incl, gb + 7), nInner = __popc(imask);
    uint32_t childBase = 0, triBase = 0;
    if (valid && g == 0) {
        if (nInner) childBase = atomicAdd(nodeTail, nInner);
        if (triTotal) triBase = atomicAdd(triTail, triTotal);
        if (childBase + nInner > regionCap) { atomicOr(p.stat, 1u); }          // cannot happen (bound in launch_wide_build); never silent
    }
    childBase = __shfl_sync(full, childBase, gb); triBase = __shfl_sync(full, triBase, gb);
    if (!valid) return;
    const bool fits = childBase + nInner <= regionCap;
    if (inner && fits) {
        WideEntry q;
        q.link = c.idx | (c.kind == K_BATCH ? kCopyFlag : 0u); q.f = c.f; q.l = c.l; q.pad = 0;
        for (int a = 0; a < 3; ++a) { q.lo[a] = c.lo[a]; q.hi[a] = c.hi[a]; }
        queue[childBase + __popc(imask & ((1u << g) - 1u))] = q;
    }
    if (!TLAS && cnt) for (uint32_t k = 0; k < cnt; ++k) triSrc[triBase + triOff + k] = (uint16_t)(c.f + k);
    // ---- child planes relative to the node origin in units of the per-axis scale, as fp16: lo rounded down, hi rounded up (the difference
    // and the power-of-two scaling are evaluated in double; both roundings go the conservative way) ----
    // extent rounded UP (the float difference may round down onto a power of two that is then smaller than the true extent: the far planes of a
    // child touching the parent's hi side would be clamped to 1.0 = below the child — found by tests/test_wide_bvh_gpu.py's enclosure check)
    const uint32_t ex = exp_byte(__double2float_ru((double)e.hi[0] - (double)e.lo[0])), ey = exp_byte(__double2float_ru((double)e.hi[1] - (double)e.lo[1])),
                   ez = exp_byte(__double2float_ru((double)e.hi[2] - (double)e.lo[2]));
    const uint32_t eb[3] = {ex, ey, ez};
    uint8_t* ob = reinterpret_cast<uint8_t*>(outNode);
    __half* oh = reinterpret_cast<__half*>(outNode);
    uint32_t meta = 0u;
    if (inner) meta = 0x20u | (24u + (uint32_t)g);
    else if (cnt) meta = (((1u << cnt) - 1u) << 5) | triOff;
    ob[24 + g] = (uint8_t)meta;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        __half hlo = __float2half_rd(0.0f), hhi = hlo;
        if (c.kind != K_EMPTY) {
            const double inv = (double)__uint_as_float((254u - eb[a]) << 23);           // 2^(127 - e), exact
            const double dlo = fmin(fmax(((double)c.lo[a] - (double)e.lo[a]) * inv, 0.0), 1.0), dhi = fmin(fmax(((double)c.hi[a] - (double)e.lo[a]) * inv, 0.0), 1.0);
            hlo = __float2half_rd(__double2float_rd(dlo)); hhi = __float2half_ru(__double2float_ru(dhi));
        }
        oh[16 + 8 * a + g] = hlo;                                                        // rows 2..4
        oh[40 + 8 * a + g] = hhi;                                                        // rows 5..7
    }
    if (g == 0) {
        outNode[0] = make_float4(e.lo[0], e.lo[1], e.lo[2], __uint_as_float(ex | (ey << 8) | (ez << 16) | (imask << 24)));
        reinterpret_cast<uint2*>(outNode)[2] = make_uint2(regionBase + childBase, triRegion + triBase);
    }
}

__device__ __forceinline__ void root_entry(const float4* __restrict__ binNodes, uint32_t n, WideEntry& e) {
    float4 rec[4];
    load_rec(binNodes, 0u, rec);
    e.link = 0u; e.f = 0u; e.l = n - 1u; e.pad = 0u;
    if (n == 1u) {                                          // src/buildBVH.cuh:31-38: the right "child" of a one-item segment is a dummy
        e.lo[0] = rec[1].x; e.lo[1] = rec[1].y; e.lo[2] = rec[1].z; e.hi[0] = rec[0].x; e.hi[1] = rec[0].y; e.hi[2] = rec[0].z;
    } else {
        e.lo[0] = fminf(rec[1].x, rec[0].w); e.lo[1] = fminf(rec[1].y, rec[1].w); e.lo[2] = fminf(rec[1].z, rec[2].w);
        e.hi[0] = fmaxf(rec[0].x, rec[2].x); e.hi[1] = fmaxf(rec[0].y, rec[2].y); e.hi[2] = fmaxf(rec[0].z, rec[2].z);
    }
}

__global__ void __launch_bounds__(kWbThreads) k_wide_blas(const WideBuildParams p) {
    __shared__ WideEntry sQueue[kWideRegion];
    __shared__ uint16_t sTriSrc[kBatchSize];
    __shared__ uint32_t sNodeTail, sTriTail;
    const uint32_t b = blockIdx.x;
    const uint32_t n = (b == p.numBatches - 1) ? (p.numTris - kBatchSize * (p.numBatches - 1)) : kBatchSize;
    const float4* binNodes = p.binNodes + (size_t)b * kBatchSize * 4;
    if (threadIdx.x == 0) { root_entry(binNodes, n, sQueue[0]); sNodeTail = 1u; sTriTail = 0u; }
    __syncthreads();
    const uint32_t group = threadIdx.x >> 3, warpGroup0 = (threadIdx.x >> 5) << 2;
    uint32_t head = 0, tail = 1;
    while (head < tail) {
        for (uint32_t base = head; base + warpGroup0 < tail; base += kWbThreads / 8) {     // same trip count for the four groups of a warp
            const uint32_t w = base + group;
            wide_process<false>(p, binNodes, sQueue, w, w < tail, &sNodeTail, &sTriTail, sTriSrc, b * (uint32_t)kWideRegion, (uint32_t)kWideRegion,
                                b * kBatchSize, n == 1u);
        }
        __syncthreads();
        head = tail; tail = min(sNodeTail, (uint32_t)kWideRegion);
        __syncthreads();
    }
    // ---- triangles in leaf-allocation order: rows 0..2 of the reference-layout Triangle, original index in row0.w ----
    const uint32_t nt = sTriTail;
    for (uint32_t j = threadIdx.x; j < nt; j += kWbThreads) {
        const uint32_t lid = __ldg(p.reorderIdx + (size_t)b * kBatchSize + sTriSrc[j]);
        const uint32_t tri = b * kBatchSize + lid;
        const float4* src = p.triRows + (size_t)tri * 8;
        float4 q1 = __ldcg(src), q2 = __ldcg(src + 1), q3 = __ldcg(src + 2);
        q1.w = __int_as_float((int)tri);
        float4* dst = p.tris + (size_t)(b * kBatchSize + j) * 3;
        dst[0] = q1; dst[1] = q2; dst[2] = q3;
    }
    if (threadIdx.x == 0) { atomicAdd(p.stat + 1, tail); if (nt != n) atomicOr(p.stat, 2u); }
}

__global__ void __launch_bounds__(kWtThreads) k_wide_tlas(const WideBuildParams p) {
    __shared__ uint32_t sNodeTail, sTriTail;
    WideEntry* queue = p.tlasQueue;
    const uint32_t B = p.numBatches;
    if (threadIdx.x == 0) {
        if (B == 1u) { WideEntry e{}; e.link = kCopyFlag | 0u; queue[0] = e; }
        else root_entry(p.binNodes, B, queue[0]);
        sNodeTail = 1u; sTriTail = 0u;
    }
    __syncthreads();
    const uint32_t group = threadIdx.x >> 3, warpGroup0 = (threadIdx.x >> 5) << 2;
    uint32_t head = 0, tail = 1;
    while (head < tail) {
        for (uint32_t base = head; base + warpGroup0 < tail; base += kWtThreads / 8) {
            const uint32_t w = base + group;
            wide_process<true>(p, p.binNodes, queue, w, w < tail, &sNodeTail, &sTriTail, nullptr, p.tlasBase, p.tlasCap, 0u, false);
        }
        __threadfence();
        __syncthreads();
        head = tail; tail = min(sNodeTail, p.tlasCap);
        __syncthreads();
    }
    if (threadIdx.x == 0) { p.stat[2] = tail; atomicAdd(p.stat + 1, tail); }
}

// Top-level leaves: node = the wide root of the batch (its children / triangles stay where they are).  Runs after BOTH collapses, so that the
// top-level tree (one CTA, ~0.1 ms of dependent steps) is collapsed CONCURRENTLY with the batches instead of behind them.
__global__ void __launch_bounds__(256) k_wide_copy(const WideBuildParams p) {
    const uint32_t w = (blockIdx.x * 256u + threadIdx.x) >> 3, g = threadIdx.x & 7u;
    if (w >= p.stat[2]) return;
    const uint32_t link = p.tlasQueue[w].link;
    if (!(link & kCopyFlag)) return;
    const float4* src = p.nodes + (size_t)((link & ~kCopyFlag) * (uint32_t)kWideRegion) * kWideStride;
    p.nodes[(size_t)(p.tlasBase + w) * kWideStride + g] = __ldcg(src + g);
}

int launch_wide_build(Context& c) {
    const uint32_t B = c.numBatches;
    const uint32_t tlasCap = 2u * B + 8u;
    const size_t nodeCount = (size_t)B * kWideRegion + tlasCap;
    if (c.wideNodeCap < nodeCount) {
        if (c.dWideNodes) cudaFree(c.dWideNodes);
        if (c.dWideTris) cudaFree(c.dWideTris);
        if (c.dWideQueue) cudaFree(c.dWideQueue);
        c.dWideNodes = nullptr; c.dWideTris = nullptr; c.dWideQueue = nullptr; c.wideNodeCap = 0;
        RTRT_CUDA(c, cudaMalloc(&c.dWideNodes, nodeCount * kWideStride * sizeof(float4)));
        RTRT_CUDA(c, cudaMalloc(&c.dWideTris, (size_t)B * kBatchSize * 3 * sizeof(float4)));
        RTRT_CUDA(c, cudaMalloc(&c.dWideQueue, (size_t)tlasCap * sizeof(WideEntry) + 64));
        c.wideNodeCap = nodeCount;
    }
    WideBuildParams p{};
    p.binNodes = reinterpret_cast<const float4*>(c.dNodes);
    p.reorderIdx = c.dReorderIdx;
    p.triRows = reinterpret_cast<const float4*>(c.dTriangles);
    p.nodes = reinterpret_cast<float4*>(c.dWideNodes); p.tris = reinterpret_cast<float4*>(c.dWideTris);
    p.numTris = c.numTris; p.numBatches = B;
    p.tlasBase = B * (uint32_t)kWideRegion; p.tlasCap = tlasCap;
    p.stat = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(c.dWideQueue) + (size_t)tlasCap * sizeof(WideEntry));
    p.tlasQueue = reinterpret_cast<WideEntry*>(c.dWideQueue);
    // wide_bvh = 2 (primary rays on the reference layout): the collapse runs on a side stream behind the build and overlaps the primary-ray
    // kernels of the frame; the first kernel that traverses the wide layout waits for it (wide_join)
    cudaStream_t st = c.stream;
    if (!c.widePrimary) {
        if (!c.wideStream) {
            RTRT_CUDA(c, cudaStreamCreateWithFlags(&c.wideStream, cudaStreamNonBlocking));
            RTRT_CUDA(c, cudaEventCreateWithFlags(&c.evWideFork, cudaEventDisableTiming));
            RTRT_CUDA(c, cudaEventCreateWithFlags(&c.evWideDone, cudaEventDisableTiming));
        }
        RTRT_CUDA(c, cudaEventRecord(c.evWideFork, c.stream));
        RTRT_CUDA(c, cudaStreamWaitEvent(c.wideStream, c.evWideFork, 0));
        st = c.wideStream;
    }
    if (!c.wideTlasStream) {
        RTRT_CUDA(c, cudaStreamCreateWithFlags(&c.wideTlasStream, cudaStreamNonBlocking));
        RTRT_CUDA(c, cudaEventCreateWithFlags(&c.evTlasFork, cudaEventDisableTiming));
        RTRT_CUDA(c, cudaEventCreateWithFlags(&c.evTlasDone, cudaEventDisableTiming));
    }
    RTRT_CUDA(c, cudaMemsetAsync(p.stat, 0, 64, st));
    RTRT_CUDA(c, cudaEventRecord(c.evTlasFork, st));
    RTRT_CUDA(c, cudaStreamWaitEvent(c.wideTlasStream, c.evTlasFork, 0));
    WideBuildParams pt = p;
    pt.binNodes = reinterpret_cast<const float4*>(c.dTlasNodes);
    k_wide_tlas<<<1, kWtThreads, 0, c.wideTlasStream>>>(pt);                     // beside the batches
    RTRT_CUDA(c, cudaEventRecord(c.evTlasDone, c.wideTlasStream));
    k_wide_blas<<<B, kWbThreads, 0, st>>>(p);
    RTRT_CUDA(c, cudaStreamWaitEvent(st, c.evTlasDone, 0));
    k_wide_copy<<<(tlasCap * 8u + 255u) / 256u, 256, 0, st>>>(p);
    if (st != c.stream) { RTRT_CUDA(c, cudaEventRecord(c.evWideDone, st)); c.widePending = true; }
    c.launches += 3;
    c.wideRoot = p.tlasBase;
    c.wideValid = true; c.wideStatPending = true;
    RTRT_CUDA(c, cudaGetLastError());
    return RTRT_OK;
}

// after a stream synchronisation: error flags of the last wide build (region overflow / triangle count mismatch)
int wide_join(Context& c) {
    if (c.widePending) { RTRT_CUDA(c, cudaStreamWaitEvent(c.stream, c.evWideDone, 0)); c.widePending = false; }
    return RTRT_OK;
}

int check_wide_error(Context& c) {
    if (!c.wideStatPending || !c.dWideQueue) return RTRT_OK;
    if (c.wideStream) RTRT_CUDA(c, cudaStreamSynchronize(c.wideStream));
    c.wideStatPending = false;
    const uint32_t tlasCap = 2u * c.numBatches + 8u;
    uint32_t st[4] = {0, 0, 0, 0};
    RTRT_CUDA(c, cudaMemcpy(st, reinterpret_cast<char*>(c.dWideQueue) + (size_t)tlasCap * sizeof(WideEntry), sizeof(st), cudaMemcpyDeviceToHost));
    c.wideNodes = st[1];
    if (st[0]) { c.wideValid = false; return c.fail(RTRT_ERR_CAPACITY, "wide BVH build: node region overflow or triangle count mismatch"); }
    return RTRT_OK;
}

}  // namespace rtrt
