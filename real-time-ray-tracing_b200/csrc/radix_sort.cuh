// Device-wide stable LSD radix sort of (key, value) pairs — the multi-CTA counterpart of block_sort_1024
// (bvh_block.cuh) for sequences that do not fit one CTA: the TLAS over more than 1024 batches and the global
// Morton pre-sort of the triangles (generalisation rule B7).  The reference has no such sort: its
// RadixSort (src/radixSort.cuh:21-246) is one block per 1024 keys, and spatial order across batches comes from an
// offline CPU sort (tool/meshProcessor.cpp:146-200).  Semantics = std::stable_sort on the key.
//
// 8-bit digits.  Per pass three launches: per-tile digit histogram -> per-digit scan over tiles -> stable scatter.
// Tile = 2048 keys per CTA (256 threads x 8 rounds); ranking inside a 32-key warp slice uses __match_any_sync.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rtrt {

constexpr int kRsThreads = 256;
#ifndef RTRT_RS_ROUNDS
#define RTRT_RS_ROUNDS 8      // keys per lane: 8 -> 60 registers, 4 CTAs/SM (C3 shuffled build 1.31 -> 1.25 ms against 16; 4: 1.27)
#endif
constexpr int kRsRounds = RTRT_RS_ROUNDS;
constexpr int kRsTile = kRsThreads * kRsRounds;      // 2048 keys per CTA

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

// Stable scatter of one tile.  Every warp owns kRsRounds x 32 consecutive keys and keeps them (one per lane and coalesced round)
// in registers; a key's rank among equal digits of its warp is the warp's running digit count before its round (a 256-entry
// shared-memory row per warp, no block barrier between rounds) + its rank inside the round (__match_any_sync).  One exchange per
// tile then turns the eight warp rows into exclusive offsets, and every key goes to
//     global base of its digit + offset of this tile inside the digit + offset of its warp inside the tile + rank inside the warp
// — through a shared-memory copy of the tile in sorted order, so that the global stores are runs of whole sectors.
// (The first version ran the whole block through histogram / barrier / scan / barrier for every 256 keys and stored straight from
// registers: 39-93 us per pass at 4 M 64-bit keys depending on the digit entropy; this one 45-60 us with 4096-key tiles, less with 2048.)
template <typename K>
__global__ void __launch_bounds__(kRsThreads) k_rs_scatter(const K* __restrict__ keysIn, const uint32_t* __restrict__ valsIn,
                                                           K* __restrict__ keysOut, uint32_t* __restrict__ valsOut, uint32_t n, int shift,
                                                           uint32_t numTiles, const uint32_t* __restrict__ tileHist,
                                                           const uint32_t* __restrict__ digitTotal) {
    constexpr int NW = kRsThreads / 32, R = kRsRounds;
    __shared__ uint32_t digitOffset[256];
    __shared__ uint16_t warpHist[NW][256];
    __shared__ uint32_t tileExcl[256];
    __shared__ uint32_t warpSum[NW];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    {
        uint32_t* z = reinterpret_cast<uint32_t*>(&warpHist[0][0]);
        for (int i = tid; i < NW * 256 / 2; i += kRsThreads) z[i] = 0u;
    }
    const uint32_t first = blockIdx.x * kRsTile + warp * (R * 32) + lane;
    K key[R]; uint32_t val[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {                                    // all loads of the tile in flight before anything depends on them
        const uint32_t i = first + r * 32;
        key[r] = i < n ? keysIn[i] : (K)0; val[r] = i < n ? valsIn[i] : 0u;
    }
    {   // global base of digit d = sum of totals of smaller digits; + this tile's exclusive offset inside the digit
        const uint32_t v = digitTotal[tid];
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t x = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += x; }
        if (lane == 31) warpSum[warp] = incl;
        __syncthreads();                                             // also publishes the zeroed warpHist
        uint32_t base = 0;
#pragma unroll
        for (int w = 0; w < NW; ++w) if (w < warp) base += warpSum[w];
        digitOffset[tid] = base + incl - v + tileHist[(size_t)tid * numTiles + blockIdx.x];
    }
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t rk[R];                                                  // rank inside the warp | digit << 16   (digit 256 = no key)
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const bool on = first + r * 32 < n;
        const uint32_t d = on ? ((uint32_t)(key[r] >> shift) & 255u) : 256u;
        const uint32_t m = __match_any_sync(0xffffffffu, d);
        const uint32_t rank = __popc(m & lt);
        const uint32_t before = on ? warpHist[warp][d] : 0u;
        __syncwarp();
        if (on && rank == 0) warpHist[warp][d] = (uint16_t)(before + __popc(m));
        __syncwarp();
        rk[r] = (before + rank) | (d << 16);
    }
    __syncthreads();
    // thread tid = digit: the warp rows become exclusive offsets inside the tile's run of that digit, the run lengths are scanned
    // into the tile-local start of every digit
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < NW; ++w) { const uint32_t c = warpHist[w][tid]; warpHist[w][tid] = (uint16_t)run; run += c; }
    {
        uint32_t incl = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t x = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += x; }
        if (lane == 31) warpSum[warp] = incl;            // (its previous readers are behind the barrier above)
        __syncthreads();
        uint32_t base = 0;
#pragma unroll
        for (int w = 0; w < NW; ++w) if (w < warp) base += warpSum[w];
        const uint32_t excl = base + incl - run;
        tileExcl[tid] = excl;
        digitOffset[tid] -= excl;                         // global position = digitOffset[d] + slot in the tile's sorted order
    }
    __syncthreads();
    // exchange through shared memory: the tile is laid out in sorted order, then written out by consecutive threads, so that a
    // digit's run (8-16 keys on average) leaves as whole sectors instead of 32 unrelated 8-byte stores per warp instruction
    extern __shared__ __align__(16) unsigned char rs_smem[];
    K* sKey = reinterpret_cast<K*>(rs_smem);
    uint32_t* sVal = reinterpret_cast<uint32_t*>(rs_smem + sizeof(K) * kRsTile);
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const uint32_t d = rk[r] >> 16;
        if (d < 256u) {
            const uint32_t slot = tileExcl[d] + warpHist[warp][d] + (rk[r] & 0xFFFFu);
            sKey[slot] = key[r]; sVal[slot] = val[r];
        }
    }
    __syncthreads();
    const uint32_t tileBase = blockIdx.x * kRsTile;
    const uint32_t cnt = n - tileBase < (uint32_t)kRsTile ? n - tileBase : (uint32_t)kRsTile;
#pragma unroll 4
    for (uint32_t slot = tid; slot < cnt; slot += kRsThreads) {
        const K k = sKey[slot];
        const uint32_t pos = digitOffset[(uint32_t)(k >> shift) & 255u] + slot;
        keysOut[pos] = k; valsOut[pos] = sVal[slot];
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
    constexpr size_t exchangeBytes = (sizeof(K) + sizeof(uint32_t)) * kRsTile;       // 24 KB (64-bit keys) / 16 KB + ~6 KB static
    // function attributes are per device and contexts may live on several GPUs of one process: set it on every call (host-side, cheap)
    if (cudaFuncSetAttribute(k_rs_scatter<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)exchangeBytes) != cudaSuccess) return -1;
    int cur = 0;
    for (int shift = beginBit; shift < endBit; shift += 8) {
        K* kin = cur ? keysAlt : keys; uint32_t* vin = cur ? valsAlt : vals;
        K* kout = cur ? keys : keysAlt; uint32_t* vout = cur ? vals : valsAlt;
        k_rs_hist<K><<<numTiles, kRsThreads, 0, stream>>>(kin, n, shift, numTiles, tileHist);
        k_rs_scan<<<256, 1024, 0, stream>>>(tileHist, numTiles, digitTotal);
        k_rs_scatter<K><<<numTiles, kRsThreads, exchangeBytes, stream>>>(kin, vin, kout, vout, n, shift, numTiles, tileHist, digitTotal);
        launches += 3;
        cur ^= 1;
    }
    return cudaGetLastError() == cudaSuccess ? cur : -1;
}

}  // namespace rtrt
