// Top level for scenes with more than 1024 batches (generalisation rule B7): the reference's
// BuildBvhLevel2 (src/bvh.cu:57-97 = UpdateTLAS + RadixSort + BuildLBVH on ONE block) caps the scene at 1023 batches
// (src/init.cu:126).  Here the same algorithm — merged root boxes, scene box, 30-bit Morton of the box centres in the
// scene box, stable sort, Karras emit, exact bottom-up refit — runs multi-CTA over global memory:
//   k_tlas_boxes -> k_tlas_keys -> device radix sort (radix_sort.cuh) -> k_tlas_karras -> k_tlas_refit
// Output arrays and their contents are those of the single-block path (and of the oracle) for any B.
#include "bvh_block.cuh"
#include "radix_sort.cuh"
#include "context.h"

namespace rtrt {

// float <-> order-preserving uint (for atomicMin / atomicMax on floats)
__device__ __forceinline__ uint32_t f2ord(float f) { const uint32_t b = __float_as_uint(f); return (b & 0x80000000u) ? ~b : (b | 0x80000000u); }
__device__ __forceinline__ float ord2f(uint32_t u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u); }

struct TlasScratch {
    Float3* centre;        // [B]
    uint32_t* sceneOrd;    // [6] min.xyz (atomicMin), max.xyz (atomicMax) as ordered uints
    uint32_t* keysAlt;     // [Bp]
    uint32_t* valsAlt;     // [Bp]
    uint32_t* parent;      // [B]  parent of internal node
    uint32_t* leafParent;  // [B]  parent of the leaf at sorted position p
    uint32_t* arrive;      // [B]
    AABB* nodeBox;         // [B]
    void* sortTmp;
};

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

size_t tlas_global_scratch_bytes(uint32_t B, uint32_t Bp) {
    return align256(sizeof(Float3) * B) + 256 + 2 * align256(4 * (size_t)Bp) + 3 * align256(4 * (size_t)B) + align256(sizeof(AABB) * (size_t)B) +
           align256(radix_sort_tmp_bytes(Bp));
}

static TlasScratch carve(void* base, uint32_t B, uint32_t Bp) {
    char* p = reinterpret_cast<char*>(base);
    TlasScratch s;
    s.centre = reinterpret_cast<Float3*>(p); p += align256(sizeof(Float3) * B);
    s.sceneOrd = reinterpret_cast<uint32_t*>(p); p += 256;
    s.keysAlt = reinterpret_cast<uint32_t*>(p); p += align256(4 * (size_t)Bp);
    s.valsAlt = reinterpret_cast<uint32_t*>(p); p += align256(4 * (size_t)Bp);
    s.parent = reinterpret_cast<uint32_t*>(p); p += align256(4 * (size_t)B);
    s.leafParent = reinterpret_cast<uint32_t*>(p); p += align256(4 * (size_t)B);
    s.arrive = reinterpret_cast<uint32_t*>(p); p += align256(4 * (size_t)B);
    s.nodeBox = reinterpret_cast<AABB*>(p); p += align256(sizeof(AABB) * (size_t)B);
    s.sortTmp = p;
    return s;
}

__global__ void k_tlas_init(uint32_t* sceneOrd, uint32_t* arrive, uint32_t B) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 3) sceneOrd[i] = 0xFFFFFFFFu; else if (i < 6) sceneOrd[i] = 0u;
    if (i < B) arrive[i] = 0u;
}

// UpdateTLAS, first half (src/updateGeometry.cuh:288-300): merged box of BLAS root k, its centre, scene box
__global__ void __launch_bounds__(256) k_tlas_boxes(const BVHNode* __restrict__ nodes, uint32_t B, AABB* __restrict__ tlasAabbs,
                                                    Float3* __restrict__ centre, uint32_t* sceneOrd) {
    const uint32_t k = blockIdx.x * 256 + threadIdx.x;
    AABB m = aabb_empty();
    if (k < B) {
        const float4* r4 = reinterpret_cast<const float4*>(nodes + (size_t)kSeg * k);
        const float4 a = __ldg(r4), b = __ldg(r4 + 1), c = __ldg(r4 + 2);
        m.max.x = sel_max(a.x, c.x); m.max.y = sel_max(a.y, c.y); m.max.z = sel_max(a.z, c.z);       // AABBCompact::GetMerged
        m.min.x = sel_min(b.x, a.w); m.min.y = sel_min(b.y, b.w); m.min.z = sel_min(b.z, c.w);
        tlasAabbs[k] = m;
        centre[k] = Float3{(m.max.x + m.min.x) / 2.0f, (m.max.y + m.min.y) / 2.0f, (m.max.z + m.min.z) / 2.0f};
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m.min.x = sel_min(m.min.x, __shfl_xor_sync(0xffffffffu, m.min.x, o)); m.min.y = sel_min(m.min.y, __shfl_xor_sync(0xffffffffu, m.min.y, o));
        m.min.z = sel_min(m.min.z, __shfl_xor_sync(0xffffffffu, m.min.z, o)); m.max.x = sel_max(m.max.x, __shfl_xor_sync(0xffffffffu, m.max.x, o));
        m.max.y = sel_max(m.max.y, __shfl_xor_sync(0xffffffffu, m.max.y, o)); m.max.z = sel_max(m.max.z, __shfl_xor_sync(0xffffffffu, m.max.z, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(sceneOrd + 0, f2ord(m.min.x)); atomicMin(sceneOrd + 1, f2ord(m.min.y)); atomicMin(sceneOrd + 2, f2ord(m.min.z));
        atomicMax(sceneOrd + 3, f2ord(m.max.x)); atomicMax(sceneOrd + 4, f2ord(m.max.y)); atomicMax(sceneOrd + 5, f2ord(m.max.z));
    }
}

// UpdateTLAS, second half (src/updateGeometry.cuh:352-362): 30-bit Morton of the centres in the scene box; pad keys = 0xFFFFFFFF
__global__ void __launch_bounds__(256) k_tlas_keys(const Float3* __restrict__ centre, const uint32_t* __restrict__ sceneOrd, uint32_t B, uint32_t Bp,
                                                   uint32_t* __restrict__ keys, uint32_t* __restrict__ ids, uint32_t* __restrict__ keysPre) {
    const uint32_t k = blockIdx.x * 256 + threadIdx.x;
    if (k >= Bp) return;
    AABB scene;
    scene.min = Float3{ord2f(sceneOrd[0]), ord2f(sceneOrd[1]), ord2f(sceneOrd[2])};
    scene.max = Float3{ord2f(sceneOrd[3]), ord2f(sceneOrd[4]), ord2f(sceneOrd[5])};
    const uint32_t key = (k < B) ? morton30(centre[k], scene) : 0xFFFFFFFFu;
    keys[k] = key; ids[k] = k;
    if (keysPre) keysPre[k] = key;
}

__device__ __forceinline__ int g_delta(const uint32_t* __restrict__ keys, int n, int i, uint32_t ki, int j) {
    if (j < 0 || j >= n) return 0;
    const uint32_t x = ki ^ __ldg(keys + j);
    return x ? __clz(x) : 32 + __clz((uint32_t)i ^ (uint32_t)j);
}

// Karras emit (src/buildBVH.cuh:60-134), one thread per internal node, links written straight into the node array
__global__ void __launch_bounds__(256) k_tlas_karras(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ ids, int n, BVHNode* __restrict__ out,
                                                     uint32_t* __restrict__ parent, uint32_t* __restrict__ leafParent) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n - 1) return;
    const uint32_t ki = __ldg(keys + i);
    const int dl = g_delta(keys, n, i, ki, i - 1), dr = g_delta(keys, n, i, ki, i + 1);
    const int d = (dr - dl) >= 0 ? 1 : -1;
    const int dmin = g_delta(keys, n, i, ki, i - d);
    int lmax = 2;
    while (g_delta(keys, n, i, ki, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (g_delta(keys, n, i, ki, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = g_delta(keys, n, i, ki, j);
    int sp = 0;
    for (int div = 2;; div <<= 1) {
        const int t = (l + div - 1) / div;
        if (g_delta(keys, n, i, ki, i + (sp + t) * d) > dnode) sp += t;
        if (t <= 1) break;
    }
    const int gamma = i + sp * d + min(d, 0);
    uint32_t L, R, lLeaf, rLeaf;
    if (min(i, j) == gamma) { L = __ldg(ids + gamma); lLeaf = 1u; leafParent[gamma] = (uint32_t)i; }
    else                    { L = (uint32_t)gamma; lLeaf = 0u; parent[gamma] = (uint32_t)i; }
    if (max(i, j) == gamma + 1) { R = __ldg(ids + gamma + 1); rLeaf = 1u; leafParent[gamma + 1] = (uint32_t)i; }
    else                        { R = (uint32_t)(gamma + 1); rLeaf = 0u; parent[gamma + 1] = (uint32_t)i; }
    reinterpret_cast<uint4*>(out + i)[3] = make_uint4(L, R, lLeaf, rLeaf);
}

__device__ __forceinline__ AABB ld_box_cg(const AABB* p) {
    const float* f = reinterpret_cast<const float*>(p);
    AABB b;
    b.min = Float3{__ldcg(f), __ldcg(f + 1), __ldcg(f + 2)}; b.max = Float3{__ldcg(f + 3), __ldcg(f + 4), __ldcg(f + 5)};
    return b;
}

// bottom-up refit (src/buildBVH.cuh:138-267): one walker per leaf; the second walker to reach a node merges its children
__global__ void __launch_bounds__(256) k_tlas_refit(int n, BVHNode* nodes, const AABB* __restrict__ leafBox, AABB* nodeBox,
                                                    const uint32_t* __restrict__ parent, const uint32_t* __restrict__ leafParent, uint32_t* arrive) {
    const int p = blockIdx.x * 256 + threadIdx.x;
    if (p >= n) return;
    uint32_t cur = leafParent[p];
    while (true) {
        __threadfence();
        if (atomicAdd(arrive + cur, 1u) == 0u) break;          // first arrival: the sibling subtree is still pending
        __threadfence();
        const uint4 lk = __ldcg(reinterpret_cast<const uint4*>(nodes + cur) + 3);
        const AABB bl = lk.z ? leafBox[lk.x] : ld_box_cg(nodeBox + lk.x);
        const AABB br = lk.w ? leafBox[lk.y] : ld_box_cg(nodeBox + lk.y);
        float4* o4 = reinterpret_cast<float4*>(nodes + cur);
        o4[0] = make_float4(bl.max.x, bl.max.y, bl.max.z, br.min.x);
        o4[1] = make_float4(bl.min.x, bl.min.y, bl.min.z, br.min.y);
        o4[2] = make_float4(br.max.x, br.max.y, br.max.z, br.min.z);
        AABB u = bl; aabb_grow(u, br);
        float* nb = reinterpret_cast<float*>(nodeBox + cur);
        __stcg(nb, u.min.x); __stcg(nb + 1, u.min.y); __stcg(nb + 2, u.min.z); __stcg(nb + 3, u.max.x); __stcg(nb + 4, u.max.y); __stcg(nb + 5, u.max.z);
        if (cur == 0u) break;
        cur = parent[cur];
    }
}

int launch_build_tlas_global(Context& c) {
    const uint32_t B = c.numBatches, Bp = c.numBatchesPadded;
    const size_t need = tlas_global_scratch_bytes(B, Bp);
    if (c.sortTmpBytes < need) {
        if (c.dSortTmp) cudaFree(c.dSortTmp);
        c.dSortTmp = nullptr; c.sortTmpBytes = 0;
        RTRT_CUDA(c, cudaMalloc(&c.dSortTmp, need));
        c.sortTmpBytes = need;
    }
    const TlasScratch s = carve(c.dSortTmp, B, Bp);
    const unsigned gB = (B + 255) / 256, gBp = (Bp + 255) / 256;
    k_tlas_init<<<gB, 256, 0, c.stream>>>(s.sceneOrd, s.arrive, B);
    k_tlas_boxes<<<gB, 256, 0, c.stream>>>(c.dNodes, B, c.dTlasAabbs, s.centre, s.sceneOrd);
    k_tlas_keys<<<gBp, 256, 0, c.stream>>>(s.centre, s.sceneOrd, B, Bp, c.dTlasMorton, c.dTlasReorderIdx, c.keepPresortMorton ? c.dTlasMortonPre : nullptr);
    c.launches += 3;
    const int where = radix_sort_pairs<uint32_t>(c.stream, c.dTlasMorton, c.dTlasReorderIdx, s.keysAlt, s.valsAlt, Bp, 0, 32, s.sortTmp, c.launches);
    if (where < 0) return c.fail(RTRT_ERR_CUDA, "TLAS radix sort launch failed");
    if (where == 1) {     // (4 passes -> result is back in the primary buffers; kept for other key widths)
        RTRT_CUDA(c, cudaMemcpyAsync(c.dTlasMorton, s.keysAlt, 4 * (size_t)Bp, cudaMemcpyDeviceToDevice, c.stream));
        RTRT_CUDA(c, cudaMemcpyAsync(c.dTlasReorderIdx, s.valsAlt, 4 * (size_t)Bp, cudaMemcpyDeviceToDevice, c.stream));
    }
    k_tlas_karras<<<gB, 256, 0, c.stream>>>(c.dTlasMorton, c.dTlasReorderIdx, (int)B, c.dTlasNodes, s.parent, s.leafParent);
    k_tlas_refit<<<gB, 256, 0, c.stream>>>((int)B, c.dTlasNodes, c.dTlasAabbs, s.nodeBox, s.parent, s.leafParent, s.arrive);
    c.launches += 2;
    RTRT_CUDA(c, cudaGetLastError());
    return RTRT_OK;
}

}  // namespace rtrt
