// Stage 1 — two-level LBVH build for sm_100a.
//
// Replaces src/bvh.cu:7-97 (BuildBvhLevel1/2 = 6 dependent launches: UpdateSceneGeometry, RadixSort,
// BuildLBVH, UpdateTLAS, RadixSort, BuildLBVH) with ONE launch when the scene has <= 1024 batches:
//   * one CTA per 1024-triangle batch does gather -> Triangle/AABB emit -> batch AABB -> Morton30 ->
//     stable radix sort -> Karras emit -> refit -> node write entirely out of shared memory, so keys,
//     boxes and links never round-trip HBM between steps;
//   * one extra CTA of the same launch builds the TLAS over the batch boxes, which every batch CTA publishes right after its
//     gather: the top level is ready before the last batches have finished their sort / emit.
// Output arrays are the reference's (src/init.cu:159-187): Triangle[ ], AABB[ ], morton[ ] (sorted),
// reorderIdx[ ], BVHNode[ ] + the TLAS equivalents, bit-identical to the reference's for scenes
// without 3+-fold duplicate keys inside a segment (see bvh_block.cuh: seg_delta).
#include "bvh_block.cuh"
#include "context.h"
#include <cstdlib>

namespace rtrt {

constexpr int kBuildThreads = 512;
#ifndef RTRT_BUILD_MINB
#define RTRT_BUILD_MINB 3          // 3 CTAs/SM: <= 40 registers/thread, 64 KB shared memory per CTA
#endif

struct BuildParams {
    const float* __restrict__ pos;       // V x 3
    const float* __restrict__ nrm;       // V x 3
    const uint32_t* __restrict__ idx;    // N x 3
    uint32_t numTris;                    // N
    uint32_t numBatches;                 // B
    Triangle* triangles; AABB* aabbs; uint32_t* morton; uint32_t* reorderIdx; BVHNode* nodes;
    uint32_t* mortonPre;                 // optional debug observation point (may be null)
    AABB* tlasAabbs; uint32_t* tlasMorton; uint32_t* tlasReorderIdx; BVHNode* tlasNodes;
    uint32_t* tlasMortonPre;             // optional
    uint32_t* ticket;                    // [0] number of batch boxes published to tlasAabbs (zeroed before launch)
                                         // [1] sticky error word: the top-level CTA gave up waiting (reported by rtrt_sync / rtrt_get_timings)
    int fuseTlas;                        // 1: one extra CTA builds the TLAS concurrently (B <= 1024)
    uint32_t spinLimit;                  // polls of the ticket before the top-level CTA gives up (option tlas_spin_limit)
    int keepAllTris;                     // option keep_all_triangles: pad items of the last batch sort behind every real triangle
};

__device__ __forceinline__ Float3 ld3(const float* __restrict__ p, uint32_t i) {
    return Float3{__ldg(p + 3 * (size_t)i), __ldg(p + 3 * (size_t)i + 1), __ldg(p + 3 * (size_t)i + 2)};
}

// ---- TLAS over B <= 1024 batch roots, executed by one CTA (src/updateGeometry.cuh:267-364 + sort + emit) ----
template <int NT>
__device__ void tlas_build_block(SegSmem& s, const BuildParams& p) {
    const int B = (int)p.numBatches;
    AABB mine = aabb_empty();
    for (int k = threadIdx.x; k < kSeg; k += NT) {
        if (k < B) {
            // box of BLAS k = AABBCompact::GetMerged of its root node (src/updateGeometry.cuh:283-300, src/geometry.h:107-110), published
            // by the CTA that builds batch k (publish_tlas_leaf below) -> L2 loads
            const float* g = reinterpret_cast<const float*>(p.tlasAabbs + k);
            AABB m;
            m.min.x = __ldcg(g); m.min.y = __ldcg(g + 1); m.min.z = __ldcg(g + 2);
            m.max.x = __ldcg(g + 3); m.max.y = __ldcg(g + 4); m.max.z = __ldcg(g + 5);
            s.leafBox[k] = m;
            s.scratch.centre[k] = Float3{(m.max.x + m.min.x) / 2.0f, (m.max.y + m.min.y) / 2.0f, (m.max.z + m.min.z) / 2.0f};
            // the reference's UpdateTLAS drops the per-thread boxes of threads 32..63 (mod 64) from its scene-box reduction
            // (src/updateGeometry.cuh:313-330 lacks the "merge tid + 32" step of :219-223): batches with k mod 256 >= 128
            // do not contribute.  Reproduced bit for bit (the Morton codes of the top level depend on it).
            if ((k & 255) < 128) aabb_grow(mine, m);
        }
    }
    const AABB scene = block_reduce_aabb<NT>(s, mine);
    for (int k = threadIdx.x; k < kSeg; k += NT) {
        const uint32_t key = (k < B) ? morton30(s.scratch.centre[k], scene) : 0xFFFFFFFFu;
        s.keys[0][k] = key; s.ids[0][k] = (uint16_t)k;
        if (p.tlasMortonPre) p.tlasMortonPre[k] = key;
    }
    __syncthreads();
    block_build_segment<NT>(s, B, p.tlasMorton, p.tlasReorderIdx, p.tlasNodes);
}

// ---- top level concurrent with the batches ----
// The TLAS needs one box per batch: GetMerged of the batch's root node = the union of its leaf boxes.  min / max are exact, so that
// union is known as soon as the batch box has been reduced (right after the gather) — long before the batch's sort / emit / node
// write are through.  Every batch CTA therefore publishes its box early (tlasAabbs[b], release, counter), and ONE EXTRA CTA of the
// same launch (blockIdx == numBatches) waits for the counter and builds the top level while the last wave of batch CTAs is still
// sorting and emitting: the serial single-CTA tail (27 us of 163 at 1 M triangles when it ran after the last batch) disappears behind
// them.  Exactness: the reference's min / max are fminf / fmaxf (bvh_block.cuh), whose result does not depend on the operand order, so
// the early reduction gives the bits the refit tree would.  A partial last batch (pad items take part in the gather box but not in
// the tree) publishes GetMerged of its root's children after the emit instead.
__device__ __forceinline__ void publish_tlas_leaf(const BuildParams& p, uint32_t b, const AABB& box) {     // one thread
    p.tlasAabbs[b] = box;
    __threadfence();
    atomicAdd(p.ticket, 1u);
}
// ---- one CTA per batch ----
template <int NT>
__global__ void __launch_bounds__(NT, RTRT_BUILD_MINB) k_build_blas(const BuildParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SegSmem& s = *reinterpret_cast<SegSmem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = NT / 32;
    const uint32_t b = blockIdx.x;
    if (b == p.numBatches) {                 // the extra CTA of the fused launch: top level, as soon as every batch box is there
        // Bounded wait: a lost publication must not hang the device.  CUDA does not promise that the batch CTAs are all resident
        // before this one, so a time-out is an ERROR, not a reason to build from stale boxes: the sticky word ticket[1] is set, the top
        // level is left untouched and the host reports RTRT_ERR_CUDA at its next synchronisation point (capi.cu: check_build_error).
        __shared__ uint32_t sSeen;
        if (tid == 0) {
            uint32_t seen = 0;
            for (uint32_t spin = 0; spin < p.spinLimit; ++spin) {
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(p.ticket) : "memory");
                if (seen >= p.numBatches) break;
                __nanosleep(100);
            }
            if (seen < p.numBatches) atomicExch(p.ticket + 1, 1u);
            sSeen = seen;
        }
        __syncthreads();
        if (sSeen < p.numBatches) return;
        tlas_build_block<NT>(s, p);
        return;
    }
    const uint32_t base = b * kSeg;
    const uint32_t count = (b == p.numBatches - 1) ? (p.numTris - kSeg * (p.numBatches - 1)) : (uint32_t)kSeg;  // src/init.cu:129-130
    const uint32_t active = (count + 3u) & ~3u;     // src/updateGeometry.cuh:93 : threads own 4 items each

    float4* tri4 = reinterpret_cast<float4*>(p.triangles + base);
    // ---- stage the batch's 3 x 1024 vertex indices in shared memory: one fully coalesced 12-KB read instead of a 4-byte load per
    // lane in front of every dependent vertex load (pad items read index 0, src/init.cu:103-115).  Lives in the upper half of
    // `scratch`, which nothing else touches before the sort (the centres occupy the lower 12 KB).
    uint32_t* sIdx = reinterpret_cast<uint32_t*>(reinterpret_cast<unsigned char*>(&s.scratch) + sizeof(Float3) * kSeg);
    {
        const size_t lim = (size_t)3 * p.numTris, off = (size_t)3 * base;
        for (uint32_t e = tid; e < 3u * kSeg; e += NT) sIdx[e] = (off + e < lim) ? __ldg(p.idx + off + e) : 0u;
    }
    __syncthreads();
    float accMin = FLT_MAX, accMax = -FLT_MAX;           // this lane's axis of the batch box
    // ---- gather + Triangle emit (src/updateGeometry.cuh:100-184).  One FLOAT per lane: three consecutive lanes read the x, y, z of
    // one vertex, so a warp-wide load touches 10 vertices (about one L1 wavefront each) instead of one component of 24 different
    // vertices three times over.  A warp step covers 8 triangles = 48 vertices = 5 loads per lane; the floats are staged in a
    // per-warp shared-memory tile of 8 x 6 float4 rows (w stays 0), from which the per-axis AABB lanes and the two fully coalesced
    // 512-byte Triangle stores of the step read.  The tile lives in keys[]/ids[] (12 KB), which nothing touches before the Morton
    // keys are written.  (Measured on B200, 1 M triangles: row-per-lane gather 0.173 ms -> 0.166 ms; keeping the next step's loads
    // in flight in registers gained nothing — 48 warps per SM already cover the latency.)
    constexpr int kTps = 8, kSteps = kSeg / (NW * kTps);
    static_assert(NW * kTps * 6 * sizeof(float4) <= sizeof(s.keys) + sizeof(s.ids), "gather tile");
    float4* stageAll = reinterpret_cast<float4*>(&s.keys[0][0]);
    float4* stageW = stageAll + warp * (kTps * 6);
    for (int e = lane; e < kTps * 6; e += 32) stageW[e] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    float* stageF = reinterpret_cast<float*>(stageW);
    const float4* stageR = stageW;
    __syncwarp();
    // Three lanes per vertex (x, y, z), ten vertices per warp-wide load; the 48 vertex slots of a step are numbered row-major
    // (slot = 8 r6 + t: r6 = 0..2 positions v1..v3, 3..5 normals n1..n3 of triangle t), so slot -> (t, r6) is a shift and a mask.
    const int vl = lane / 3, vc = lane - 3 * vl;
    auto issue = [&](int st, float (&v)[5]) {
        const uint32_t i0 = (uint32_t)(st * NW + warp) * kTps;
#pragma unroll
        for (int j = 0; j < 5; ++j) {
            const int vs = 10 * j + vl, t = vs & 7, r6 = vs >> 3;
            v[j] = 0.0f;
            if (vl < 10 && vs < kTps * 6) {     // items past `active` read index 0 like the pad items (sIdx holds 0 there)
                const uint32_t vid = sIdx[3u * (i0 + t) + (r6 >= 3 ? r6 - 3 : r6)];
                v[j] = __ldg((r6 < 3 ? p.pos : p.nrm) + (3u * vid + vc));
            }
        }
    };
    float cur[5];
#pragma unroll 2
    for (int st = 0; st < kSteps; ++st) {
        issue(st, cur);
        const uint32_t i0 = (uint32_t)(st * NW + warp) * kTps;
#pragma unroll
        for (int j = 0; j < 5; ++j) {
            const int vs = 10 * j + vl, t = vs & 7, r6 = vs >> 3;
            if (vl < 10 && vs < kTps * 6) stageF[(t * 6 + r6) * 4 + vc] = cur[j];
        }
        __syncwarp();
        {   // per-axis AABB / centre: lane = 4 t + axis
            const int t = lane >> 2, ax = lane & 3;
            const uint32_t i = i0 + t;
            if (ax < 3 && i < active) {
                const float a = stageF[(t * 6 + 0) * 4 + ax], b = stageF[(t * 6 + 1) * 4 + ax], c = stageF[(t * 6 + 2) * 4 + ax];
                const float mn = sel_min(a, sel_min(b, c));                                       // src/updateGeometry.cuh:172-177
                float mx = sel_max(a, sel_max(b, c));
                const float eps = 1.1920928955078125e-07f;                                        // MachineEpsilon()
                mx = mn + sel_max(mx - mn, eps * mx);
                float* lb = reinterpret_cast<float*>(&s.leafBox[i]);                              // AABB = {min.xyz, max.xyz}
                lb[ax] = mn; lb[3 + ax] = mx;
                reinterpret_cast<float*>(&s.scratch.centre[i])[ax] = ((a + b) + c) / 3.0f;       // :180
                accMin = sel_min(accMin, mn); accMax = sel_max(accMax, mx);
            }
        }
#pragma unroll
        for (int k = 0; k < 2; ++k) {       // Triangle emit: rows 0..2 = v1..v3, 3 = v4 (unused), 4..6 = n1..n3, 7 = n4 (unused)   :131-137
            const int q = 32 * k + lane, t = q >> 3, r = q & 7;
            const uint32_t i = i0 + t;
            if (i < active) tri4[(size_t)i * 8 + r] = (r & 3) == 3 ? make_float4(0.0f, 0.0f, 0.0f, 0.0f) : stageR[t * 6 + (r < 4 ? r : r - 1)];
        }
        __syncwarp();
    }
    const int sub = lane & 3;               // the axis this lane accumulated
    AABB mine = aabb_empty();
    if (sub == 0) { mine.min.x = accMin; mine.max.x = accMax; }
    else if (sub == 1) { mine.min.y = accMin; mine.max.y = accMax; }
    else if (sub == 2) { mine.min.z = accMin; mine.max.z = accMax; }
    const AABB batchBox = block_reduce_aabb<NT>(s, mine);     // barriers inside also publish leafBox/centre
    const bool tlasEarly = p.fuseTlas && count == (uint32_t)kSeg;          // block-uniform
    if (tlasEarly && tid == 0) publish_tlas_leaf(p, b, batchBox);

    // ---- per-item AABB array (24 KB contiguous per batch) + Morton keys ----
    if (active == (uint32_t)kSeg) {                          // 24 KB = 1536 float4, three per thread
        const float4* lb = reinterpret_cast<const float4*>(s.leafBox);
        float4* ga = reinterpret_cast<float4*>(p.aabbs + base);
        for (int e = tid; e < kSeg * 6 / 4; e += NT) ga[e] = lb[e];
    } else {
        const float* lb = reinterpret_cast<const float*>(s.leafBox);
        float* ga = reinterpret_cast<float*>(p.aabbs + base);
        for (uint32_t e = tid; e < active * 6u; e += NT) ga[e] = lb[e];
    }
    for (int i = tid; i < kSeg; i += NT) {
        // The reference gives the (up to 3) pad items of a partial last batch real keys while the hierarchy only takes the first `count`
        // sorted entries: pad items (degenerate, at vertex 0) can push real triangles out of the tree (src/init.cu:103-115,
        // src/updateGeometry.cuh:93).  Reproduced by default; keep_all_triangles sorts the pad items last instead.
        const uint32_t keyed = p.keepAllTris ? count : active;
        const uint32_t key = ((uint32_t)i < keyed) ? morton30(s.scratch.centre[i], batchBox) : 0xFFFFFFFFu;  // src/kernel.cu:279
        s.keys[0][i] = key; s.ids[0][i] = (uint16_t)i;
        if (p.mortonPre) p.mortonPre[base + i] = key;
    }
    __syncthreads();

    block_sort_publish<NT>(s, p.morton + base, p.reorderIdx + base);
    block_emit_write<NT>(s, (int)count, p.nodes + base);
    if (p.fuseTlas && !tlasEarly && tid == 0) {                  // partial last batch: GetMerged of the root node's two child boxes
        AABB bl, br;
        if (count == 1) { bl = s.leafBox[0]; br.min = Float3{0.0f, 0.0f, 0.0f}; br.max = br.min; }      // src/buildBVH.cuh:31-38
        else { bl = child_box(s, unpack_link(s.u.link.childL[0])); br = child_box(s, unpack_link(s.u.link.childR[0])); }
        AABB m;
        m.max.x = sel_max(bl.max.x, br.max.x); m.max.y = sel_max(bl.max.y, br.max.y); m.max.z = sel_max(bl.max.z, br.max.z);
        m.min.x = sel_min(bl.min.x, br.min.x); m.min.y = sel_min(bl.min.y, br.min.y); m.min.z = sel_min(bl.min.z, br.min.z);
        publish_tlas_leaf(p, b, m);
    }
}

// ------------------------------------------------------------------------------------------------
int launch_build_bvh(Context& c) {
    if (c.numTris == 0) return c.fail(RTRT_ERR_STATE, "rtrt_build_bvh: no mesh set");
    { int rc = wait_geometry(c); if (rc) return rc; }       // a pending rtrt_update_vertices upload (copy stream)
    if (c.globalSort) { int rc = launch_presort(c); if (rc) return rc; }
    BuildParams p{};
    p.pos = c.dPos; p.nrm = c.dNrm; p.idx = c.globalSort ? c.dIdxSorted : c.dIdx; p.numTris = c.numTris; p.numBatches = c.numBatches;
    p.triangles = c.dTriangles; p.aabbs = c.dAabbs; p.morton = c.dMorton; p.reorderIdx = c.dReorderIdx; p.nodes = c.dNodes;
    p.mortonPre = c.keepPresortMorton ? c.dMortonPre : nullptr;
    p.tlasAabbs = c.dTlasAabbs; p.tlasMorton = c.dTlasMorton; p.tlasReorderIdx = c.dTlasReorderIdx; p.tlasNodes = c.dTlasNodes;
    p.tlasMortonPre = c.keepPresortMorton ? c.dTlasMortonPre : nullptr;
    p.ticket = c.dTicket;
    p.fuseTlas = (c.numBatches <= (uint32_t)kSeg) ? 1 : 0;
    p.spinLimit = c.tlasSpinLimit; p.keepAllTris = c.keepAllTris ? 1 : 0;

    const size_t smem = sizeof(SegSmem);
    if (!c.buildAttrSet) {                           // function attributes are per device: the flag lives in the (per-device) context
        RTRT_CUDA(c, cudaFuncSetAttribute(k_build_blas<kBuildThreads>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        // 3 x (64 KB + 1 KB reserved) = 195 KB: ask for the 196 KB carve-out (85 % of 228 KB rounds up to it), the rest of the SM's 256 KB is L1
        const char* cv = getenv("RTRT_BUILD_CARVEOUT");
        RTRT_CUDA(c, cudaFuncSetAttribute(k_build_blas<kBuildThreads>, cudaFuncAttributePreferredSharedMemoryCarveout, cv ? atoi(cv) : 85));
        c.buildAttrSet = true;
    }
    RTRT_CUDA(c, cudaMemsetAsync(c.dTicket, 0, sizeof(uint32_t), c.stream));
    if (p.fuseTlas) c.buildErrPending = true;
    k_build_blas<kBuildThreads><<<c.numBatches + (p.fuseTlas ? 1u : 0u), kBuildThreads, smem, c.stream>>>(p);
    c.launches += 1;
    { int rc = geometry_consumed(c); if (rc) return rc; }   // the vertex buffers may be overwritten from here on
    if (!p.fuseTlas) {
        int rc = launch_build_tlas_global(c);     // B > 1024: multi-CTA top level (bvh_tlas_global.cu)
        if (rc) return rc;
    }
    RTRT_CUDA(c, cudaGetLastError());
    return RTRT_OK;
}

}  // namespace rtrt
