// Device-wide stable LSD radix sort of (key, value) pairs — the multi-CTA counterpart of block_sort_1024
// (bvh_block.cuh) for sequences that do not fit one CTA: the TLAS over more than 1024 batches and the global
// Morton pre-sort of the triangles (generalisation rule B7).  The reference has no such sort: its
// RadixSort (src/radixSort.cuh:21-246) is one block per 1024 keys, and spatial order across batches comes from an
// offline CPU sort (tool/meshProcessor.cpp:146-200).  Semantics = std::stable_sort on the key.
//
// 8-bit digits.  Per pass three launches: per-tile digit histogram -> per-digit scan over tiles -> stable scatter.
// Tile = 4096 keys per CTA (256 threads x 16 rounds); ranking inside a 32-key warp slice uses __match_any_sync.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rtrt {

constexpr int kRsThreads = 256;
constexpr int kRsRounds = 16;
constexpr int kRsTile = kRsThreads * kRsRounds;      // 4096 keys per CTA

template <typename K>
__global__ void __launch_bounds__(kRsThreads) k_rs_hist(const K* __restrict__ keys, uint32_t n, int shift, uint32_t numTiles,
                                                        uint32_t* __restrict__ tileHist /* [256][numTiles] */) {
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0u;
    __syncthreads();
    const uint32_t base = blockIdx.x * kRsTile;
#pragma unroll 4
    for (int r = 0; r < kRsRounds; ++r) {
        const uint32_t i = base + r * kRsThreads + threadIdx.x;
        if (i < n) atomicAdd(&h[(uint32_t)(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    tileHist[(size_t)threadIdx.x * numTiles + blockIdx.x] = h[threadIdx.x];
}

// one CTA per digit: exclusive scan of that digit's counts over the tiles (in place) + the digit total
static __global__ void __launch_bounds__(1024) k_rs_scan(uint32_t* __restrict__ tileHist, uint32_t numTiles, uint32_t* __restrict__ digitTotal) {
    __shared__ uint32_t warpSum[32];
    __shared__ uint32_t carry;
    uint32_t* row = tileHist + (size_t)blockIdx.x * numTiles;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0u;
    __syncthreads();
    for (uint32_t t0 = 0; t0 < numTiles; t0 += 1024) {
        const uint32_t t = t0 + threadIdx.x;
        const uint32_t v = t < numTiles ? row[t] : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t x = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += x; }
        if (lane == 31) warpSum[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = warpSum[lane], wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t x = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += x; }
            warpSum[lane] = wi - w;                       // exclusive warp offsets
        }
        __syncthreads();
        const uint32_t excl = carry + warpSum[warp] + incl - v;
        if (t < numTiles) row[t] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) digitTotal[blockIdx.x] = carry;
}

template <typename K>
__global__ void __launch_bounds__(kRsThreads) k_rs_scatter(const K* __restrict__ keysIn, const uint32_t* __restrict__ valsIn,
                                                           K* __restrict__ keysOut, uint32_t* __restrict__ valsOut, uint32_t n, int shift,
                                                           uint32_t numTiles, const uint32_t* __restrict__ tileHist,
                                                           const uint32_t* __restrict__ digitTotal) {
    constexpr int NW = kRsThreads / 32;
    __shared__ uint32_t digitOffset[256];          // running output position of the next key with this digit
    __shared__ uint16_t warpCount[NW][256];
    __shared__ uint32_t warpSum[NW];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // global base of digit d = sum of totals of smaller digits; + this tile's exclusive offset inside the digit
    {
        const uint32_t v = digitTotal[tid];
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t x = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += x; }
        if (lane == 31) warpSum[warp] = incl;
        __syncthreads();
        uint32_t base = 0;
#pragma unroll
        for (int w = 0; w < NW; ++w) if (w < warp) base += warpSum[w];
        digitOffset[tid] = base + incl - v + tileHist[(size_t)tid * numTiles + blockIdx.x];
    }
    const uint32_t lt = (1u << lane) - 1u;
    const uint32_t tileBase = blockIdx.x * kRsTile;
    __shared__ uint32_t warpBase[NW][256];
#pragma unroll 1
    for (int r = 0; r < kRsRounds; ++r) {
        uint32_t* wc32 = reinterpret_cast<uint32_t*>(&warpCount[0][0]);
        for (int i = tid; i < NW * 256 / 2; i += kRsThreads) wc32[i] = 0u;
        __syncthreads();                                            // also publishes digitOffset of the previous round
        const uint32_t i = tileBase + r * kRsThreads + tid;
        const bool on = i < n;
        K key = 0; uint32_t val = 0, d = 256u;                      // 256 = "no key": its own match group, never counted
        if (on) { key = keysIn[i]; val = valsIn[i]; d = (uint32_t)(key >> shift) & 255u; }
        const uint32_t m = __match_any_sync(0xffffffffu, d);
        const uint32_t rank = __popc(m & lt);
        if (on && rank == 0) warpCount[warp][d] = (uint16_t)__popc(m);
        __syncthreads();
        {   // thread tid = digit: exclusive prefix over the warps (stable order = round, warp, lane), advance the running offset
            uint32_t run = digitOffset[tid];
#pragma unroll
            for (int w = 0; w < NW; ++w) { warpBase[w][tid] = run; run += warpCount[w][tid]; }
            digitOffset[tid] = run;
        }
        __syncthreads();
        if (on) {
            const uint32_t pos = warpBase[warp][d] + rank;
            keysOut[pos] = key; valsOut[pos] = val;
        }
    }
}

// Scratch needed by radix_sort_pairs for n keys.
inline size_t radix_sort_tmp_bytes(uint32_t n) {
    const size_t numTiles = (n + kRsTile - 1) / kRsTile;
    return (256 * numTiles + 256) * sizeof(uint32_t);
}

// Sorts by key bits [beginBit, endBit) (multiples of 8).  Ping-pongs between (keys, vals) and (keysAlt, valsAlt);
// returns 0 if the result is in (keys, vals), 1 if it is in the Alt buffers, negative on a launch error.
template <typename K>
inline int radix_sort_pairs(cudaStream_t stream, K* keys, uint32_t* vals, K* keysAlt, uint32_t* valsAlt, uint32_t n,
                            int beginBit, int endBit, void* tmp, unsigned long long& launches) {
    if (n == 0) return 0;
    const uint32_t numTiles = (n + kRsTile - 1) / kRsTile;
    uint32_t* tileHist = reinterpret_cast<uint32_t*>(tmp);
    uint32_t* digitTotal = tileHist + (size_t)256 * numTiles;
    int cur = 0;
    for (int shift = beginBit; shift < endBit; shift += 8) {
        K* kin = cur ? keysAlt : keys; uint32_t* vin = cur ? valsAlt : vals;
        K* kout = cur ? keys : keysAlt; uint32_t* vout = cur ? vals : valsAlt;
        k_rs_hist<K><<<numTiles, kRsThreads, 0, stream>>>(kin, n, shift, numTiles, tileHist);
        k_rs_scan<<<256, 1024, 0, stream>>>(tileHist, numTiles, digitTotal);
        k_rs_scatter<K><<<numTiles, kRsThreads, 0, stream>>>(kin, vin, kout, vout, n, shift, numTiles, tileHist, digitTotal);
        launches += 3;
        cur ^= 1;
    }
    return cudaGetLastError() == cudaSuccess ? cur : -1;
}

}  // namespace rtrt
