// Global Morton pre-sort of the triangles on the device (generalisation rule B7).
//
// The reference's GPU sort is per 1024-triangle batch; spatial coherence ACROSS batches comes from an offline CPU tool
// (tool/meshProcessor.cpp:35-62,146-200: 60-bit Morton code of the centroid = 20 bits per axis in the mesh bounding box,
// std::sort, triangles rewritten in that order).  For dynamic or unsorted geometry (BASELINE config 3) that step has to
// run per frame on the GPU: option "global_sort" = 1 makes rtrt_build_bvh run
//     k_ps_bounds (vertex bounding box) -> k_ps_codes (60-bit code per triangle; 63-bit = 21 bits per axis with option
//     "morton_bits" = 63) -> device radix sort of (code, id),
//     8 x 8-bit passes (radix_sort.cuh) -> k_ps_gather (index buffer in sorted order)
// in front of the batch build.  Sorted position -> original triangle id is observable as RTRT_BUF_TRI_PERMUTATION.
// Semantics = stable sort of the ids by code (equal codes keep input order).
#include "bvh_block.cuh"
#include "radix_sort.cuh"
#include "context.h"

namespace rtrt {

__device__ __forceinline__ uint32_t f2ord_ps(float f) { const uint32_t b = __float_as_uint(f); return (b & 0x80000000u) ? ~b : (b | 0x80000000u); }
__device__ __forceinline__ float ord2f_ps(uint32_t u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u); }

__global__ void k_ps_init(uint32_t* ord) { if (threadIdx.x < 3) ord[threadIdx.x] = 0xFFFFFFFFu; else if (threadIdx.x < 6) ord[threadIdx.x] = 0u; }

__global__ void __launch_bounds__(256) k_ps_bounds(const float* __restrict__ pos, uint32_t V, uint32_t* ord) {
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t v = blockIdx.x * 256 + threadIdx.x; v < V; v += gridDim.x * 256) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { const float p = __ldg(pos + 3 * (size_t)v + a); mn[a] = fminf(mn[a], p); mx[a] = fmaxf(mx[a], p); }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { mn[a] = fminf(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o)); mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o)); }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { atomicMin(ord + a, f2ord_ps(mn[a])); atomicMax(ord + 3 + a, f2ord_ps(mx[a])); }
    }
}

// tool/meshProcessor.cpp:35-62 : q = (uint)clamp((c - lo) / ext * (2^20 - 1)), code = M30(q >> 10) << 30 | M30(q & 1023)
__device__ __forceinline__ uint32_t quant20(float c, float lo, float hi) {
    float ext = hi - lo;
    if (!(ext > 0.0f)) ext = 1.0f;
    const float scale = 1048575.0f;
    float u = (c - lo) / ext * scale;
    u = fminf(fmaxf(u, 0.0f), scale);
    return __float2uint_rz(u);
}

// 63-bit mode (option "morton_bits" = 63; generalisation rule B7): 21 bits per axis, one interleave x | y << 1 | z << 2.
// Same quantisation rule as the tool's with 2^21 - 1 in place of 2^20 - 1 (float32: 24 mantissa bits cover 21).
__device__ __forceinline__ uint32_t quant21(float c, float lo, float hi) {
    float ext = hi - lo;
    if (!(ext > 0.0f)) ext = 1.0f;
    const float scale = 2097151.0f;
    float u = (c - lo) / ext * scale;
    u = fminf(fmaxf(u, 0.0f), scale);
    return __float2uint_rz(u);
}
__device__ __forceinline__ unsigned long long expand21(uint32_t v) {        // 21 bits -> every third bit of 63
    unsigned long long x = v & 0x1FFFFFull;
    x = (x | (x << 32)) & 0x001F00000000FFFFull;
    x = (x | (x << 16)) & 0x001F0000FF0000FFull;
    x = (x | (x << 8)) & 0x100F00F00F00F00Full;
    x = (x | (x << 4)) & 0x10C30C30C30C30C3ull;
    x = (x | (x << 2)) & 0x1249249249249249ull;
    return x;
}

__global__ void __launch_bounds__(256) k_ps_codes(const float* __restrict__ pos, const uint32_t* __restrict__ idx, uint32_t N, const uint32_t* __restrict__ ord,
                                                  unsigned long long* __restrict__ keys, uint32_t* __restrict__ vals, int bits63) {
    const uint32_t t = blockIdx.x * 256 + threadIdx.x;
    if (t >= N) return;
    const uint32_t i0 = __ldg(idx + 3 * (size_t)t), i1 = __ldg(idx + 3 * (size_t)t + 1), i2 = __ldg(idx + 3 * (size_t)t + 2);
    uint32_t q[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float c = ((__ldg(pos + 3 * (size_t)i0 + a) + __ldg(pos + 3 * (size_t)i1 + a)) + __ldg(pos + 3 * (size_t)i2 + a)) / 3.0f;
        q[a] = bits63 ? quant21(c, ord2f_ps(ord[a]), ord2f_ps(ord[3 + a])) : quant20(c, ord2f_ps(ord[a]), ord2f_ps(ord[3 + a]));
    }
    if (bits63) {
        keys[t] = expand21(q[0]) | (expand21(q[1]) << 1) | (expand21(q[2]) << 2);
        vals[t] = t;
        return;
    }
    const unsigned long long lo30 = expand10(q[0] & 1023u) | (expand10(q[1] & 1023u) << 1) | (expand10(q[2] & 1023u) << 2);
    const unsigned long long hi30 = expand10(q[0] >> 10) | (expand10(q[1] >> 10) << 1) | (expand10(q[2] >> 10) << 2);
    keys[t] = (hi30 << 30) | lo30;
    vals[t] = t;
}

__global__ void __launch_bounds__(256) k_ps_gather(const uint32_t* __restrict__ idx, const uint32_t* __restrict__ perm, uint32_t N,
                                                   uint32_t* __restrict__ idxSorted, uint32_t* __restrict__ permOut) {
    const uint32_t t = blockIdx.x * 256 + threadIdx.x;
    if (t >= N) return;
    const uint32_t s = __ldg(perm + t);
    idxSorted[3 * (size_t)t] = __ldg(idx + 3 * (size_t)s);
    idxSorted[3 * (size_t)t + 1] = __ldg(idx + 3 * (size_t)s + 1);
    idxSorted[3 * (size_t)t + 2] = __ldg(idx + 3 * (size_t)s + 2);
    if (permOut != perm) permOut[t] = s;
}

static size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }

int launch_presort(Context& c) {
    const uint32_t N = c.numTris;
    const size_t need = 256 + 2 * al256(8 * (size_t)N) + al256(4 * (size_t)N) + al256(radix_sort_tmp_bytes(N));
    if (c.presortBytes < need) {
        if (c.dPresort) cudaFree(c.dPresort);
        c.dPresort = nullptr; c.presortBytes = 0;
        RTRT_CUDA(c, cudaMalloc(&c.dPresort, need));
        c.presortBytes = need;
    }
    if (!c.dIdxSorted) RTRT_CUDA(c, cudaMalloc((void**)&c.dIdxSorted, 12 * (size_t)N));
    if (!c.dTriPerm) RTRT_CUDA(c, cudaMalloc((void**)&c.dTriPerm, 4 * (size_t)N));
    char* p = reinterpret_cast<char*>(c.dPresort);
    uint32_t* ord = reinterpret_cast<uint32_t*>(p); p += 256;
    unsigned long long* keys = reinterpret_cast<unsigned long long*>(p); p += al256(8 * (size_t)N);
    unsigned long long* keysAlt = reinterpret_cast<unsigned long long*>(p); p += al256(8 * (size_t)N);
    uint32_t* valsAlt = reinterpret_cast<uint32_t*>(p); p += al256(4 * (size_t)N);
    void* tmp = p;
    const int smCount = c.smCount;
    k_ps_init<<<1, 32, 0, c.stream>>>(ord);
    k_ps_bounds<<<smCount * 8, 256, 0, c.stream>>>(c.dPos, c.numVerts, ord);
    k_ps_codes<<<(N + 255) / 256, 256, 0, c.stream>>>(c.dPos, c.dIdx, N, ord, keys, c.dTriPerm, c.mortonBits == 63 ? 1 : 0);
    c.launches += 3;
    const int where = radix_sort_pairs<unsigned long long>(c.stream, keys, c.dTriPerm, keysAlt, valsAlt, N, 0, 64, tmp, c.launches);
    if (where < 0) return c.fail(RTRT_ERR_CUDA, "global Morton sort launch failed");
    const uint32_t* perm = where ? valsAlt : c.dTriPerm;
    k_ps_gather<<<(N + 255) / 256, 256, 0, c.stream>>>(c.dIdx, perm, N, c.dIdxSorted, c.dTriPerm);
    c.launches += 1;
    RTRT_CUDA(c, cudaGetLastError());
    return RTRT_OK;
}

}  // namespace rtrt
