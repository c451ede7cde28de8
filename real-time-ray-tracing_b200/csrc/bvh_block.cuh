// Block-level building blocks of the LBVH build: every step of one 1024-item segment runs out of
// shared memory inside ONE CTA (the reference round-trips HBM between three kernels per level:
// src/bvh.cu:12,29,44).  Shared by the BLAS kernel (one CTA per 1024-triangle batch) and the TLAS
// kernel (one CTA over <= 1024 batch roots).
//
// All arithmetic that decides output bits (AABB padding, Morton scaling, min/max selection) is
// written with the reference's operation order and compiled with --fmad=false.
#pragma once
#include "rtrt_types.h"
#include <cuda_runtime.h>
#include <float.h>


namespace rtrt {

constexpr int kSeg = 1024;                 // items per segment (BatchSize)
constexpr uint32_t kLeafBit = 0x80000000u; // child link flag inside shared memory

// The reference's scalar min(a, b) / max(a, b) calls (min3f / max3f src/linearMath.h:463-464, AABBCompact::GetMerged src/geometry.h:107-110,
// ...) do NOT reach its own select-style templates (src/linearMath.h:25-31): for two floats, overload resolution prefers CUDA's
// non-template ::min(float, float) / ::max(float, float), i.e. fminf / fmaxf — one FMNMX, NaN loses, -0 < +0.  The result is therefore
// independent of the operand order down to the sign of a zero (pinned against the reference's own kernels by
// tests/test_refcuda_gpu.py::test_signed_zero_boxes_bitwise), which is what lets the batch boxes be reduced in any order.
__device__ __forceinline__ float sel_min(float a, float b) { return fminf(a, b); }
__device__ __forceinline__ float sel_max(float a, float b) { return fmaxf(a, b); }

__device__ __forceinline__ AABB aabb_empty() {
    AABB b; b.min = Float3{FLT_MAX, FLT_MAX, FLT_MAX}; b.max = Float3{-FLT_MAX, -FLT_MAX, -FLT_MAX}; return b;
}
__device__ __forceinline__ void aabb_grow(AABB& a, const AABB& b) {
    a.min.x = sel_min(a.min.x, b.min.x); a.min.y = sel_min(a.min.y, b.min.y); a.min.z = sel_min(a.min.z, b.min.z);
    a.max.x = sel_max(a.max.x, b.max.x); a.max.y = sel_max(a.max.y, b.max.y); a.max.z = sel_max(a.max.z, b.max.z);
}

// 10 bits per axis -> 30-bit code, x lowest (src/updateGeometry.cuh:13-27)
__device__ __forceinline__ uint32_t expand10(uint32_t x) {
    x = (x | (x << 16)) & 0x030000FFu;
    x = (x | (x << 8)) & 0x0300F00Fu;
    x = (x | (x << 4)) & 0x030C30C3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}
// src/updateGeometry.cuh:254-258 — cvt.rzi.u32.f32 saturates and maps NaN to 0 (flat axis gives 0/0)
__device__ __forceinline__ uint32_t morton30(Float3 c, const AABB& box) {
    float ux = (c.x - box.min.x) / (box.max.x - box.min.x);
    float uy = (c.y - box.min.y) / (box.max.y - box.min.y);
    float uz = (c.z - box.min.z) / (box.max.z - box.min.z);
    return expand10(__float2uint_rz(ux * 1023.0f)) | (expand10(__float2uint_rz(uy * 1023.0f)) << 1) |
           (expand10(__float2uint_rz(uz * 1023.0f)) << 2);
}

// child links in shared memory: 16 bits (index < 1024, leaf flag in bit 15) <-> the 32-bit form used in registers (kLeafBit)
__device__ __forceinline__ uint16_t pack_link(uint32_t l) { return (uint16_t)((l & 0x7FFFu) | ((l >> 16) & 0x8000u)); }
__device__ __forceinline__ uint32_t unpack_link(uint16_t v) { return ((uint32_t)v & 0x7FFFu) | (((uint32_t)v & 0x8000u) << 16); }

// Shared-memory working set of one segment (64 KB: three CTAs and their 1 KB system reservations fit the 196 KB carve-out, which leaves
// 60 KB of the SM's 256 KB to L1 for the vertex gather instead of 28 KB).  `scratch` is time-multiplexed:
//   gather phase : item centres (12 KB) ; sort phase : digit histograms (16 KB) ; refit : node boxes (24 KB)
struct SegSmem {
    AABB leafBox[kSeg];                        // 24 KB  per-item boxes, indexed by original (pre-sort) id
    union {
        Float3   centre[kSeg];
        uint16_t hist[32][256];
        AABB     nodeBox[kSeg];
    } scratch;                                 // 24 KB
    uint32_t keys[2][kSeg];                    //  8 KB  ping-pong; after the sort keys[1] holds the rendezvous slots of the emit
    uint16_t ids[2][kSeg];                     //  4 KB  ping-pong
    // 4 KB, time-multiplexed: block reduction / sort bookkeeping (dead before the emit) | child links of the emit (11 bits + leaf flag)
    union {
        struct { uint16_t childL[kSeg], childR[kSeg]; } link;
        struct { uint16_t digitBase[256]; uint32_t warpSum[8]; float red[6][16]; AABB segBox; uint32_t misc[4]; } aux;
    } u;
};

// ---- block-wide AABB reduction (exact; min/max are order independent) ----
template <int NT>
__device__ __forceinline__ AABB block_reduce_aabb(SegSmem& s, AABB v) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        v.min.x = sel_min(v.min.x, __shfl_xor_sync(0xffffffffu, v.min.x, o));
        v.min.y = sel_min(v.min.y, __shfl_xor_sync(0xffffffffu, v.min.y, o));
        v.min.z = sel_min(v.min.z, __shfl_xor_sync(0xffffffffu, v.min.z, o));
        v.max.x = sel_max(v.max.x, __shfl_xor_sync(0xffffffffu, v.max.x, o));
        v.max.y = sel_max(v.max.y, __shfl_xor_sync(0xffffffffu, v.max.y, o));
        v.max.z = sel_max(v.max.z, __shfl_xor_sync(0xffffffffu, v.max.z, o));
    }
    if (lane == 0) {
        s.u.aux.red[0][warp] = v.min.x; s.u.aux.red[1][warp] = v.min.y; s.u.aux.red[2][warp] = v.min.z;
        s.u.aux.red[3][warp] = v.max.x; s.u.aux.red[4][warp] = v.max.y; s.u.aux.red[5][warp] = v.max.z;
    }
    __syncthreads();
    if (warp == 0) {
        constexpr int NW = NT / 32;
        AABB w = aabb_empty();
        if (lane < NW) {
            w.min = Float3{s.u.aux.red[0][lane], s.u.aux.red[1][lane], s.u.aux.red[2][lane]};
            w.max = Float3{s.u.aux.red[3][lane], s.u.aux.red[4][lane], s.u.aux.red[5][lane]};
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            w.min.x = sel_min(w.min.x, __shfl_xor_sync(0xffffffffu, w.min.x, o));
            w.min.y = sel_min(w.min.y, __shfl_xor_sync(0xffffffffu, w.min.y, o));
            w.min.z = sel_min(w.min.z, __shfl_xor_sync(0xffffffffu, w.min.z, o));
            w.max.x = sel_max(w.max.x, __shfl_xor_sync(0xffffffffu, w.max.x, o));
            w.max.y = sel_max(w.max.y, __shfl_xor_sync(0xffffffffu, w.max.y, o));
            w.max.z = sel_max(w.max.z, __shfl_xor_sync(0xffffffffu, w.max.z, o));
        }
        if (lane == 0) s.u.aux.segBox = w;
    }
    __syncthreads();
    return s.u.aux.segBox;
}

// ---- stable LSD radix sort of the 1024 (key,id) pairs in s.keys[0]/s.ids[0]; 4 passes x 8 bits ----
// Replaces src/radixSort.cuh:21-246 (8 passes x 4 bits, ballot emulation of match-any, ~80 barriers).
// Ranking inside a 32-key chunk uses __match_any_sync; chunk histograms are combined by a per-digit
// column scan (one thread per digit) + one 256-wide block scan.  Result lands back in keys[0]/ids[0].
template <int NT>
__device__ __forceinline__ void block_sort_1024(SegSmem& s) {
    static_assert(NT >= 256 && NT <= 512 && NT % 32 == 0 && kSeg % NT == 0, "block size");
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = NT / 32;
    constexpr int CHUNKS_PER_WARP = 32 / NW;            // 32 chunks of 32 keys
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll 1
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = pass * 8;
        const uint32_t* ksrc = s.keys[pass & 1]; uint32_t* kdst = s.keys[(pass + 1) & 1];
        const uint16_t* isrc = s.ids[pass & 1];  uint16_t* idst = s.ids[(pass + 1) & 1];
        uint32_t* h32 = reinterpret_cast<uint32_t*>(&s.scratch.hist[0][0]);
        for (int i = tid; i < 32 * 256 / 2; i += NT) h32[i] = 0u;
        __syncthreads();
        uint32_t key[CHUNKS_PER_WARP], rank[CHUNKS_PER_WARP];
#pragma unroll
        for (int r = 0; r < CHUNKS_PER_WARP; ++r) {
            const int c = warp * CHUNKS_PER_WARP + r;
            key[r] = ksrc[c * 32 + lane];
            const uint32_t d = (key[r] >> shift) & 255u;
            const uint32_t m = __match_any_sync(0xffffffffu, d);
            rank[r] = __popc(m & lt);
            if (rank[r] == 0) s.scratch.hist[c][d] = (uint16_t)__popc(m);
        }
        __syncthreads();
        // column scan over the 32 chunks for digit d = tid, then exclusive scan of the 256 totals
        uint32_t total = 0;
        if (tid < 256) {
#pragma unroll 8
            for (int c = 0; c < 32; ++c) {
                const uint32_t v = s.scratch.hist[c][tid];
                s.scratch.hist[c][tid] = (uint16_t)total;
                total += v;
            }
        }
        uint32_t incl = total;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += n;
        }
        if (tid < 256 && lane == 31) s.u.aux.warpSum[warp] = incl;
        __syncthreads();
        if (tid < 256) {
            uint32_t base = 0;
#pragma unroll
            for (int w = 0; w < 8; ++w) if (w < warp) base += s.u.aux.warpSum[w];
            s.u.aux.digitBase[tid] = (uint16_t)(base + incl - total);
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < CHUNKS_PER_WARP; ++r) {
            const int c = warp * CHUNKS_PER_WARP + r;
            const uint32_t d = (key[r] >> shift) & 255u;
            const uint32_t pos = (uint32_t)s.u.aux.digitBase[d] + s.scratch.hist[c][d] + rank[r];
            kdst[pos] = key[r];
            idst[pos] = isrc[c * 32 + lane];
        }
        __syncthreads();
    }
}

// delta(i,j) of Karras 2012 with the index tie-break for equal keys (the reference's LCP,
// src/buildBVH.cuh:8-14, has none and emits broken hierarchies when a key occurs 3+ times).
__device__ __forceinline__ int seg_delta(const uint32_t* keys, int n, int i, uint32_t ki, int j) {
    if (j < 0 || j >= n) return 0;
    const uint32_t x = ki ^ keys[j];
    return x ? __clz(x) : 32 + __clz((uint32_t)i ^ (uint32_t)j);
}

// ---- hierarchy emit + refit in ONE bottom-up pass (replaces src/buildBVH.cuh:60-134 emit + :138-267 refit) ----
// The binary radix tree over the sorted keys (Karras 2012, what the reference's top-down BuildLBVH emits) is unique, so it can be
// grown from the leaves (Apetrei 2014): a finished range [f,l] joins the neighbour it shares the longer prefix with —
//   delta(f-1,f) > delta(l,l+1)  -> it is the RIGHT child of the node that splits at gamma = f-1,
//   otherwise                     -> it is the LEFT  child of the node that splits at gamma = l
// (the two are never equal for distinct keys, and seg_delta's index tie-break makes equal keys distinct).  The first child to
// reach a split leaves its outer bound there; the second completes the range, and with it everything Karras' numbering needs:
// a left child is numbered by the END of its range, a right child by its START, the root is 0; the children of the node that
// splits at gamma are therefore gamma and gamma+1 (or the leaf items there).  Same topology, same indices, same links as the
// top-down emit — and the node box is merged on the way up, so there is no separate refit walk and no parent table.
// One walker per leaf, ~2 key comparisons per level instead of ~30 binary-search probes per node.
__device__ __forceinline__ int seg_bound(const uint32_t* keys, int j) {                  // delta(j, j+1)
    const uint32_t x = keys[j] ^ keys[j + 1];
    return x ? __clz(x) : 32 + __clz((uint32_t)j ^ (uint32_t)(j + 1));
}
__device__ __forceinline__ bool seg_joins_left(const uint32_t* keys, int n, int f, int l) {
    return l == n - 1 || (f > 0 && seg_bound(keys, f - 1) > seg_bound(keys, l));
}

// Shared-memory accessors of the emit walk.  The boxes written by one walker are read by the walker that arrives second at the
// same split; the two are ordered by the acq_rel exchange on the split's slot (release of the first arrival's stores, acquire
// before the second arrival's loads), so plain (weak) 8-byte accesses are enough — no block-wide fence on the critical path.
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ AABB lds_box(const AABB* p) {
    AABB b; const uint32_t a = smem_addr(p);
    asm volatile("ld.shared.v2.f32 {%0,%1}, [%6];\n\tld.shared.v2.f32 {%2,%3}, [%6+8];\n\tld.shared.v2.f32 {%4,%5}, [%6+16];"
                 : "=f"(b.min.x), "=f"(b.min.y), "=f"(b.min.z), "=f"(b.max.x), "=f"(b.max.y), "=f"(b.max.z) : "r"(a) : "memory");
    return b;
}
__device__ __forceinline__ void sts_box(AABB* p, const AABB& b) {
    const uint32_t a = smem_addr(p);
    asm volatile("st.shared.v2.f32 [%6], {%0,%1};\n\tst.shared.v2.f32 [%6+8], {%2,%3};\n\tst.shared.v2.f32 [%6+16], {%4,%5};"
                 :: "f"(b.min.x), "f"(b.min.y), "f"(b.min.z), "f"(b.max.x), "f"(b.max.y), "f"(b.max.z), "r"(a) : "memory");
}
__device__ __forceinline__ uint32_t atoms_exch_acq_rel(uint32_t* p, uint32_t v) {
    uint32_t old;
    asm volatile("atom.acq_rel.cta.shared.exch.b32 %0, [%1], %2;" : "=r"(old) : "r"(smem_addr(p)), "r"(v) : "memory");
    return old;
}
__device__ __forceinline__ AABB child_box(SegSmem& s, uint32_t link) {
    return lds_box((link & kLeafBit) ? &s.leafBox[link & ~kLeafBit] : &s.scratch.nodeBox[link]);
}

// One walker per leaf.  A walker carries its subtree's range [f,l], link and box in registers together with the two prefix lengths
// that decide where it joins next: dF = delta(f-1,f) (-1 at the segment start), dL = delta(l,l+1) (-2 at the segment end).  The
// message left in a split's slot is the first arrival's OUTER bound and the delta beyond it, so the second arrival extends its range
// without touching the keys again; of the two child boxes only the sibling's is loaded.
// Measured on B200 (1 M triangles, whole build): walkers that re-derive both deltas from the keys, load both child boxes and fence
// around a relaxed exchange 0.191 ms -> this form 0.173 ms.  Tried and dropped: the completing walker writing its 64-byte node
// straight to global memory instead of block_write_nodes (+3 %: the release fence then waits on the global stores), a relaxed
// exchange without the fence (-1.5 %, but not ordered by the PTX memory model), byte-sized chunk counts in the sort that the
// column scan clears as it reads (one barrier less per pass, no change), an L2 prefetch of a later batch's indices (no change).
template <int NT>
__device__ __forceinline__ void block_emit_refit(SegSmem& s, int n) {
    const uint32_t* keys = s.keys[0];
    uint32_t* slot = s.keys[1];                // dead after block_sort_1024 (the sorted keys live in keys[0])
    constexpr uint32_t kEmpty = 0xFFFFFFFFu;
    for (int i = threadIdx.x; i < n - 1; i += NT) slot[i] = kEmpty;
    __syncthreads();
    for (int p = threadIdx.x; p < n; p += NT) {
        int f = p, l = p;
        int dF = p > 0 ? seg_bound(keys, p - 1) : -1;
        int dL = p < n - 1 ? seg_bound(keys, p) : -2;
        uint32_t myLink = (uint32_t)s.ids[0][p] | kLeafBit;
        AABB mine = s.leafBox[myLink & ~kLeafBit];
        while (true) {
            const bool left = dF > dL;                                         // joins the left neighbour = RIGHT child of the split f-1
            const int gamma = left ? f - 1 : l;
            const uint32_t msg = (uint32_t)(left ? l : f) | ((uint32_t)((left ? dL : dF) + 2) << 16);
            const uint32_t other = atoms_exch_acq_rel(&slot[gamma], msg);
            if (other == kEmpty) break;                                        // first arrival: the sibling subtree is pending
            const int ob = (int)(other & 0xFFFFu), od = (int)(other >> 16) - 2;
            const int sibLeafPos = left ? gamma : gamma + 1;                   // the sibling is a leaf iff its range is this one item
            const uint32_t sibLink = (ob == sibLeafPos) ? ((uint32_t)s.ids[0][sibLeafPos] | kLeafBit) : (uint32_t)sibLeafPos;
            const AABB sib = child_box(s, sibLink);
            if (left) { f = ob; dF = od; } else { l = ob; dL = od; }
            const uint32_t L = left ? sibLink : myLink, R = left ? myLink : sibLink;
            AABB bl, br;                                                       // child order matters to sel_min/sel_max on signed zeros
            bl.min.x = left ? sib.min.x : mine.min.x; bl.min.y = left ? sib.min.y : mine.min.y; bl.min.z = left ? sib.min.z : mine.min.z;
            bl.max.x = left ? sib.max.x : mine.max.x; bl.max.y = left ? sib.max.y : mine.max.y; bl.max.z = left ? sib.max.z : mine.max.z;
            br.min.x = left ? mine.min.x : sib.min.x; br.min.y = left ? mine.min.y : sib.min.y; br.min.z = left ? mine.min.z : sib.min.z;
            br.max.x = left ? mine.max.x : sib.max.x; br.max.y = left ? mine.max.y : sib.max.y; br.max.z = left ? mine.max.z : sib.max.z;
            // Karras numbering: a right child carries the START of its range, a left child the END, the root (dF = -1 > dL = -2) is 0
            const bool root = f == 0 && l == n - 1;
            const int idx = (dF > dL) ? f : l;
            s.u.link.childL[idx] = pack_link(L); s.u.link.childR[idx] = pack_link(R);
            mine = bl; aabb_grow(mine, br);
            if (root) break;
            sts_box(&s.scratch.nodeBox[idx], mine);
            myLink = (uint32_t)idx;
        }
    }
    __syncthreads();
}

// ---- write the n-1 nodes of the segment: 4 lanes per 64-byte node, one float4 each (coalesced) ----
template <int NT>
__device__ __forceinline__ void block_write_nodes(const SegSmem& s, int n, BVHNode* __restrict__ out) {
    float4* o4 = reinterpret_cast<float4*>(out);
    for (int e = threadIdx.x; e < (n - 1) * 4; e += NT) {
        const int i = e >> 2, q = e & 3;
        const uint32_t L = unpack_link(s.u.link.childL[i]), R = unpack_link(s.u.link.childR[i]);
        float4 v;
        if (q == 3) {
            v.x = __uint_as_float(L & ~kLeafBit); v.y = __uint_as_float(R & ~kLeafBit);
            v.z = __uint_as_float(L >> 31);       v.w = __uint_as_float(R >> 31);
        } else {
            const AABB* bl = (L & kLeafBit) ? &s.leafBox[L & ~kLeafBit] : &s.scratch.nodeBox[L];
            const AABB* br = (R & kLeafBit) ? &s.leafBox[R & ~kLeafBit] : &s.scratch.nodeBox[R];
            if (q == 0)      v = make_float4(bl->max.x, bl->max.y, bl->max.z, br->min.x);
            else if (q == 1) v = make_float4(bl->min.x, bl->min.y, bl->min.z, br->min.y);
            else             v = make_float4(br->max.x, br->max.y, br->max.z, br->min.z);
        }
        o4[e] = v;
    }
}

// n == 1: src/buildBVH.cuh:31-38 — left box = the only item, right box = the point AABB(0,0)
__device__ __forceinline__ void write_single_leaf_node(const AABB& b, BVHNode* out) {
    BVHNode nd;
    nd.box1maxX = b.max.x; nd.box1maxY = b.max.y; nd.box1maxZ = b.max.z;
    nd.box1minX = b.min.x; nd.box1minY = b.min.y; nd.box1minZ = b.min.z;
    nd.box2maxX = nd.box2maxY = nd.box2maxZ = 0.0f; nd.box2minX = nd.box2minY = nd.box2minZ = 0.0f;
    nd.idxLeft = 0; nd.idxRight = 0; nd.isLeftLeaf = 1; nd.isRightLeaf = 1;
    *out = nd;
}

// sort -> publish sorted keys/reorderIdx | Karras -> refit -> nodes, for a segment whose keys/ids and leafBox are already in shared
// memory (two halves: the batch kernel publishes its top-level box in between).
template <int NT>
__device__ __forceinline__ void block_sort_publish(SegSmem& s, uint32_t* __restrict__ gKeys, uint32_t* __restrict__ gReorder) {
    block_sort_1024<NT>(s);
    for (int i = threadIdx.x; i < kSeg; i += NT) { gKeys[i] = s.keys[0][i]; gReorder[i] = s.ids[0][i]; }
}
template <int NT>
__device__ __forceinline__ void block_emit_write(SegSmem& s, int n, BVHNode* __restrict__ gNodes) {
    if (n == 1) {
        if (threadIdx.x == 0) write_single_leaf_node(s.leafBox[0], gNodes);
        return;
    }
    block_emit_refit<NT>(s, n);
    block_write_nodes<NT>(s, n, gNodes);
}
template <int NT>
__device__ __forceinline__ void block_build_segment(SegSmem& s, int n, uint32_t* __restrict__ gKeys,
                                                    uint32_t* __restrict__ gReorder, BVHNode* __restrict__ gNodes) {
    block_sort_publish<NT>(s, gKeys, gReorder);
    block_emit_write<NT>(s, n, gNodes);
}

}  // namespace rtrt
