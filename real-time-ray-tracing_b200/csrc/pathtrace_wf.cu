// Stage 2, wavefront driver (default): the reference's PathTrace megakernel (src/pathtrace.cuh:11-128) re-cut for a
// GPU without RT cores.  Every path runs the SAME per-pixel sequence of operations as the megakernel (same blue-noise
// samples, same guards, same five segments), but the work is split where the rays are:
//
//   k_pt_primary   one thread per pixel: sampler, thin-lens ray, primary TraverseBvh (coherent) -> hit record (16 B/px).
//   k_pt_logic_first  one thread per pixel: segment-0 outputs (depth, motion, mask) and the path logic up to the point
//                  where the next ray is needed; that ray + 80 B of path state go to HBM and the pixel id is appended
//                  (warp-aggregated) to the ray queue.  Sky pixels finish here.
//   k_pt_trace     persistent CTAs; each lane pulls a queued ray, traverses it with the shared trav_step(), writes
//                  (t, tri, u, v) and immediately REFILLS from the queue when enough lanes of its warp have gone idle,
//                  so the wildly varying lengths of incoherent bounce rays do not leave lanes empty.  Finished rays
//                  are binned into the shade queue: hits from the front, misses from the back.
//   k_pt_logic     compacted: one thread per queued path applies the hit (normal, material, ray cone), runs the
//                  glossy / diffuse interactions of its segment(s) and either queues the next ray or writes the pixel.
//   (trace, logic) x 4 rounds = at most 5 rays per path like the reference; empty rounds cost one idle launch.
//
// Path state in HBM (per pixel, SoA float4): orig|coneWidth, dir|coneSpread, beta0|albedo.x, beta1|albedo.y,
// albedo.z|lightIdx|flags|mask+seg  and the hit record t|ua|va|tri = 96 B/px, touched only for live paths.
#include "pathtrace_common.cuh"
#include "wide_bvh.cuh"

namespace rtrt {

#ifndef RTRT_TRACE_THREADS
#define RTRT_TRACE_THREADS 128
#endif
#ifndef RTRT_LOGIC_THREADS
#define RTRT_LOGIC_THREADS 128
#endif
constexpr int kTraceThreads = RTRT_TRACE_THREADS;
constexpr int kLogicThreads = RTRT_LOGIC_THREADS;
#ifndef RTRT_PRIMARY_MINB
#define RTRT_PRIMARY_MINB 14       // <= 72 registers for the wide layout's primary kernel (96 unbounded): 0.670 -> 0.644 ms (12: 0.655); the reference layout's needs 64 anyway
#endif
#ifndef RTRT_TRACE_MINB
#define RTRT_TRACE_MINB 6        // measured: 2.80 -> 2.73 ms per path-traced 1080p frame against no bound (4, 5: no change; 8: 2.75)
#endif
#ifndef RTRT_LOGIC_MINB
#define RTRT_LOGIC_MINB 5      // <= 102 registers, 5 CTAs/SM: measured 2.83 -> 2.71 ms per path-traced 1080p frame against ptxas' own choice (128 registers, 4 CTAs/SM); 6 is the same, 3 no change
#endif
#ifndef RTRT_REFILL_IDLE
#define RTRT_REFILL_IDLE 12
#endif
constexpr int kRefillIdle = RTRT_REFILL_IDLE;   // refill a warp once this many of its lanes are idle

struct WfBuffers {
    float4* st0; float4* st1; float4* st2; float4* st3; float4* st4; float4* hit;
    uint32_t* rayQueue;      // pixel ids whose next ray is to be traced (written by logic, read by trace)
    uint32_t* shadeQueue;    // the same ids after tracing, hits packed from the front, misses from the back (written by trace)
    uint32_t capacity;       // entries per queue (= pixels)
    uint32_t* counters;      // [r] rays queued for round r (0..3)   [8+r] fetch cursor of k_pt_trace
                             // [16+r] hits of round r               [24+r] misses of round r
};

enum : uint32_t { F_HITLIGHT = 1u, F_PROCESSED = 2u, F_OCCLUDED = 4u, F_SHADOW = 8u, F_DIFFUSE = 16u, F_HIT = 32u };

VD void save_state(const WfBuffers& B, uint32_t pix, const RayState& r, int seg, unsigned short mask) {
    B.st0[pix] = make_float4(r.orig.x, r.orig.y, r.orig.z, r.rayConeWidth);
    B.st1[pix] = make_float4(r.dir.x, r.dir.y, r.dir.z, r.rayConeSpread);
    B.st2[pix] = make_float4(r.beta0.x, r.beta0.y, r.beta0.z, r.albedo.x);
    B.st3[pix] = make_float4(r.beta1.x, r.beta1.y, r.beta1.z, r.albedo.y);
    const uint32_t flags = (r.hitLight ? F_HITLIGHT : 0u) | (r.isHitProcessed ? F_PROCESSED : 0u) | (r.isOccluded ? F_OCCLUDED : 0u) |
                           (r.isShadowRay ? F_SHADOW : 0u) | (r.isDiffuse ? F_DIFFUSE : 0u) | (r.hit ? F_HIT : 0u) |
                           ((uint32_t)r.matType << 8);
    B.st4[pix] = make_float4(r.albedo.z, __int_as_float(r.lightIdx), __uint_as_float(flags), __uint_as_float((uint32_t)mask | ((uint32_t)seg << 16)));
}

VD void load_state(const WfBuffers& B, uint32_t pix, RayState& r, int& seg, unsigned short& mask) {
    const float4 a = B.st0[pix], b = B.st1[pix], c = B.st2[pix], d = B.st3[pix], e = B.st4[pix];
    r.orig = Float3{a.x, a.y, a.z}; r.rayConeWidth = a.w;
    r.dir = Float3{b.x, b.y, b.z}; r.rayConeSpread = b.w;
    r.beta0 = Float3{c.x, c.y, c.z}; r.beta1 = Float3{d.x, d.y, d.z};
    r.albedo = Float3{c.w, d.w, e.x};
    r.lightIdx = __float_as_int(e.y);
    const uint32_t flags = __float_as_uint(e.z), ms = __float_as_uint(e.w);
    r.hitLight = flags & F_HITLIGHT; r.isHitProcessed = flags & F_PROCESSED; r.isOccluded = flags & F_OCCLUDED;
    r.isShadowRay = flags & F_SHADOW; r.isDiffuse = flags & F_DIFFUSE; r.hit = flags & F_HIT;
    r.matType = (int)((flags >> 8) & 0xFFu);
    mask = (unsigned short)(ms & 0xFFFFu); seg = (int)(ms >> 16);
    r.matId = 0; r.objectIdx = -1; r.normal = f3(0, -1, 0); r.fakeNormal = f3(0.0f); r.pos = f3(0.0f);
    r.depth = 0.0f; r.offset = 0.0f; r.normalDotRayDir = 0.0f; r.isRayIntoSurface = false;
}

VD Float4 blue_rand4(const PathTraceParams& P, int ix, int iy, int k) {                  // src/pathtrace.cuh:53-62
    const int si = P.frameNum * 4 + k;
    return Float4{blue_rand(P.bn, ix, iy, si, 0), blue_rand(P.bn, ix, iy, si, 1), blue_rand(P.bn, ix, iy, si, 2), blue_rand(P.bn, ix, iy, si, 3)};
}

VD Float3 center_ray_dir(const Camera& cam, int ix, int iy) {                            // the centre ray of GenerateRay (src/raygen.cuh:7-38)
    Float2 uvc = (Float2{(float)ix, (float)iy} + 0.5f) * cam.inversedResolution;
    uvc = uvc * -2.0f + 1.0f;
    return normalize(cam.adjustedFront + cam.adjustedLeft * uvc.x + cam.adjustedUp * uvc.y);
}

// The interactions that follow RaySceneIntersect number `seg` (src/pathtrace.cuh:84-105), then advance through the
// segments whose intersect is a guarded no-op.  Returns true when the path needs its next ray traced (state to be saved
// with the returned seg), false when the path is complete (seg == 5).
VD bool advance_path(const PathTraceParams& P, RayState& r, int& seg, int ix, int iy, ushort4* normalOut) {
    while (true) {
        if (seg < 4) {
            const bool glossyActs = !(r.hitLight || r.isDiffuse || r.isOccluded);
            float g = 0.0f;
            if (glossyActs) g = blue_rand(P.bn, ix, iy, P.frameNum * 4, seg);             // rnd[0].{x,y,z,w} for seg 0..3
            glossy_interaction(r, g);
        }
        if (seg == 2 || seg == 3) {
            const bool first = seg == 2;
            if (!(r.hitLight || !r.isDiffuse || r.isOccluded)) {
                if (first) r.centerRaydir = center_ray_dir(P.camera, ix, iy);
                Float3 beta = first ? r.beta1 : r.beta0;
                const Float4 ra = blue_rand4(P, ix, iy, first ? 0 : 2), rb = blue_rand4(P, ix, iy, first ? 1 : 3);
                diffuse_interaction(first ? 0 : 1, P, r, beta, ra, rb);
                if (first) r.beta1 = beta; else r.beta0 = beta;
            }
            if (first) {
                Float3 n = r.fakeNormal;
                if (isnan3(n)) n = f3(0.0f);
                normalOut[(size_t)iy * P.width + ix] = make_ushort4(f2h_bits(n.x), f2h_bits(n.y), f2h_bits(n.z), f2h_bits(0.0f));
            }
        }
        ++seg;
        if (seg >= 5) return false;
        if (needs_ray(r)) return true;
    }
}

VD void finalize_path(const PathTraceParams& P, const RayState& r, unsigned short mask, int ix, int iy) {
    Float3 L0 = f3(0.0f);                                                               // GetLightSource src/light.cuh:275-305
    if (r.hitLight && !r.isOccluded && r.matType == MAT_SKY) L0 = env_light(P.env, r.dir);
    Float3 L2 = (L0 * r.beta0) * r.beta1;
    if (isnan3(L2)) L2 = f3(0.0f);                                                      // NAN_DETECTER src/pathtrace.cuh:113-117
    L2 = Float3{clampf(L2.x, 0.0f, 10.0f), clampf(L2.y, 0.0f, 10.0f), clampf(L2.z, 0.0f, 10.0f)};
    L2 = L2 / r.albedo;
    const size_t i = (size_t)iy * P.width + ix;
    P.color[i] = make_ushort4(f2h_bits(L2.x), f2h_bits(L2.y), f2h_bits(L2.z), mask);
    P.albedo[i] = make_ushort4(f2h_bits(r.albedo.x), f2h_bits(r.albedo.y), f2h_bits(r.albedo.z), f2h_bits(0.0f));
}

// warp-aggregated append of this lane's pixel to a queue; must be reached by every lane of the warp
VD void queue_append(bool emit, uint32_t entry, uint32_t* queue, uint32_t* counter, int lane) {
    const unsigned m = __ballot_sync(0xffffffffu, emit);
    if (m == 0u) return;
    const int leader = __ffs(m) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(counter, (uint32_t)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (emit) queue[base + __popc(m & ((1u << lane) - 1u))] = entry;
}

// Ray-queue entry = pixel id | kAnyHitBit.  The bit marks rays for which only "does the ray reach the sky" matters, so the traversal
// may stop at the first accepted triangle instead of completing the closest-hit search (the reference always completes it,
// src/traverse.cuh:37-48, and classifies afterwards):
//   * shadow rays: any geometry in the way sets isOccluded (the sampled light is always the environment, kEnvLightId), which
//     turns every later segment into a no-op and the radiance into 0;
//   * every ray of the last segment (intersect #4): nothing follows but GetLightSource, which is 0 unless the ray missed.
// The traversal is the reference's up to the first acceptance, so "hit anything" is decided identically (iteration cap and
// dropped pushes included) and all frame outputs are unchanged; only the hit record of such a ray is not the closest one.
constexpr uint32_t kAnyHitBit = 0x80000000u;

VD AABB scene_box(const SceneView& sc) {                     // merged box of TLAS node 0 (src/traverse.h:127-129)
    const float4 a = __ldg(sc.tlasRows), b = __ldg(sc.tlasRows + 1), c = __ldg(sc.tlasRows + 2);
    AABB box;
    box.max = Float3{sel_max(a.x, c.x), sel_max(a.y, c.y), sel_max(a.z, c.z)};
    box.min = Float3{sel_min(b.x, a.w), sel_min(b.y, b.w), sel_min(b.z, c.w)};
    return box;
}

// ---------------------------------------------------------------------------------------------------------------------
// k_pt_primary: per pixel — sampler, thin-lens camera ray, primary traversal -> hit record.  Nothing else: the kernel
// stays at traversal-kernel register counts, and all shading happens compacted in k_pt_logic.
// ---------------------------------------------------------------------------------------------------------------------
VD void init_path(const PathTraceParams& P, RayState& r, Float2& sampleUv, int ix, int iy) {
    r.beta0 = f3(1.0f); r.beta1 = f3(1.0f);
    r.hitLight = false; r.lightIdx = kDefaultLightId; r.isHitProcessed = false; r.isOccluded = false; r.isShadowRay = false;
    r.isDiffuse = false; r.hit = false; r.matId = 0; r.matType = 0; r.objectIdx = -1;
    r.normal = f3(0, -1, 0); r.fakeNormal = f3(0.0f); r.albedo = f3(1.0f); r.pos = f3(0.0f);
    r.rayConeWidth = 0.0f; r.depth = 0.0f; r.offset = 0.0f; r.normalDotRayDir = 0.0f; r.isRayIntoSurface = false;
    r.rayConeSpread = ray_cone_width(P.camera, ix, iy);
    const Float4 rnd0 = blue_rand4(P, ix, iy, 0);
    generate_ray(r.orig, r.dir, r.centerRaydir, sampleUv, P.camera, ix, iy, Float2{rnd0.x, rnd0.y}, Float2{rnd0.z, rnd0.w});
}

template <bool COUNT, class Trav>
__global__ void __launch_bounds__(kPtThreads, RTRT_PRIMARY_MINB) k_pt_primary(const PathTraceParams P, const WfBuffers B) {
    __shared__ uint2 sRaw[16 * kPtThreads];
    const int tid = threadIdx.y * kPtBlockX + threadIdx.x, lane = tid & 31;
    const int ix = blockIdx.x * kPtBlockX + threadIdx.x;
    const int jc = blockIdx.y * kPtBlockY + threadIdx.y, jy = chunk_row(P, jc), iy = frame_row(P, jy);
    uint32_t nodeVisits = 0, leafTests = 0;
    {
        const bool inside = ix < P.width && jc < P.chunkRows && jy < P.localRows;
        Trav tr;
        uint2 spill[Trav::kSpill];
        tr.setup(P.sc, sRaw, tid, kPtThreads, spill);
        if (inside) {
            const Float4 rnd0 = blue_rand4(P, ix, iy, 0);
            Float3 orig, dir, centerDir; Float2 sampleUv;
            generate_ray(orig, dir, centerDir, sampleUv, P.camera, ix, iy, Float2{rnd0.x, rnd0.y}, Float2{rnd0.z, rnd0.w});
            tr.begin(P.sc, orig, dir);
        }
        if constexpr (Trav::kWide) {
            bool active = inside;                                 // every lane of the warp steps together (the wide layout votes inside)
            while (__any_sync(0xffffffffu, active)) {
                if (tr.template step_warp<COUNT>(P.sc, 0, active)) active = false;
            }
        } else {
            if (inside) { while (!tr.template step<COUNT>(P.sc, 0)) {} }        // the reference's loop, one ray per thread
        }
        if (inside) {
            if (COUNT) { nodeVisits = tr.nodeVisits(); leafTests = tr.leafTests(); }
            B.hit[(size_t)iy * P.width + ix] = make_float4(tr.t(), tr.ua(), tr.va(), __int_as_float(tr.tri()));
        }
    }
    if (COUNT && P.rayCounter) {
        for (int o = 16; o > 0; o >>= 1) { nodeVisits += __shfl_xor_sync(0xffffffffu, nodeVisits, o); leafTests += __shfl_xor_sync(0xffffffffu, leafTests, o); }
        if (lane == 0) { atomicAdd(P.rayCounter + 1, (unsigned long long)nodeVisits); atomicAdd(P.rayCounter + 2, (unsigned long long)leafTests); }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// k_pt_trace: persistent wavefront traversal with lane refill.  Queue entries are pixel ids; rays come from st0/st1.
// Trav = ParityTrav (the reference's node / triangle arrays, its loop step by step) or WideTrav (8-wide compressed layout).
// ---------------------------------------------------------------------------------------------------------------------
template <bool COUNT, class Trav>
__global__ void __launch_bounds__(kTraceThreads, RTRT_TRACE_MINB) k_pt_trace(const SceneView sc, const WfBuffers B, const uint32_t* __restrict__ countPtr,
                                                            uint32_t* fetchPtr, uint32_t* hitCount, uint32_t* missCount, unsigned long long* visitCounters) {
    const uint32_t* __restrict__ queue = B.rayQueue;
    __shared__ uint2 sRaw[16 * kTraceThreads];
    const int tid = threadIdx.x, lane = tid & 31;
    const uint32_t n = *countPtr;
    if (n == 0u) return;
    // Small queues (the later bounce rounds: a few 10^4 rays) are spread thin: with fewer rays than lanes on the device every warp
    // takes only `perWarp` rays at a time, so a warp's time is the longest of a few traversals (and each of its steps waits on a
    // few node loads) instead of the longest of 32.
    const uint32_t totalWarps = gridDim.x * (kTraceThreads / 32);
    const uint32_t perWarp = min(32u, max(1u, (n + totalWarps - 1u) / totalWarps));
    const unsigned laneMask = perWarp >= 32u ? 0xffffffffu : ((1u << perWarp) - 1u);
    const int refillIdle = perWarp >= 32u ? kRefillIdle : 1;
    bool active = false, exhausted = false;
    uint32_t pix = 0;
    int anyHit = 0;
    Trav tr;
    uint2 spill[Trav::kSpill];
    uint32_t nodeVisits = 0, leafTests = 0;
    tr.setup(sc, sRaw, tid, kTraceThreads, spill);
    while (true) {
        // ---- refill idle lanes from the queue ----
        const unsigned idle = __ballot_sync(0xffffffffu, !active) & laneMask;
        if (!exhausted && idle != 0u) {
            const int leader = __ffs(idle) - 1;
            uint32_t base = 0;
            if (lane == leader) base = atomicAdd(fetchPtr, (uint32_t)__popc(idle));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (!active && ((idle >> lane) & 1u)) {
                const uint32_t q = base + __popc(idle & ((1u << lane) - 1u));
                if (q < n) {
                    const uint32_t entry = __ldg(queue + q);
                    pix = entry & ~kAnyHitBit; anyHit = (int)(entry >> 31);
                    const float4 o = B.st0[pix], d = B.st1[pix];
                    tr.begin(sc, Float3{o.x, o.y, o.z}, Float3{d.x, d.y, d.z});
                    active = true;
                }
            }
            exhausted = base + (uint32_t)__popc(idle) >= n;
        }
        if (!__any_sync(0xffffffffu, active)) break;
        // ---- traverse until enough lanes have finished to make a refill worthwhile ----
        while (true) {
            bool finishedNow = false;
            auto finish = [&]() {
                B.hit[pix] = make_float4(tr.t(), tr.ua(), tr.va(), __int_as_float(tr.tri()));
                if (COUNT) {
                    nodeVisits += tr.nodeVisits(); leafTests += tr.leafTests();
                    atomicMax(fetchPtr + 32, tr.iterations()); atomicAdd(fetchPtr + 40, tr.iterations());      // diagnostics: [40+r] max, [48+r] sum of iterations
                    atomicAdd(fetchPtr + 48, tr.nodeVisits()); atomicAdd(fetchPtr + 52, tr.leafTests());      // [56+r] node visits, [60+r] leaf tests of the round
                }
                active = false; finishedNow = true;
            };
            if constexpr (Trav::kWide) {                       // every lane of the warp steps together (the wide layout votes inside)
                if (tr.template step_warp<COUNT>(sc, anyHit, active)) finish();
            } else {
                if (active) { if (tr.template step<COUNT>(sc, anyHit)) finish(); }      // this nesting, not a flag tested after the step: 1.34 -> 1.22 ms (72 vs 78 registers)
            }
            // shade queue: hits from the front, misses from the back, so that k_pt_logic warps are all-surface or all-sky
            const unsigned fin = __ballot_sync(0xffffffffu, finishedNow);
            if (fin) {
                const unsigned mh = __ballot_sync(0xffffffffu, finishedNow && tr.tri() >= 0), mm = fin & ~mh;
                uint32_t bh = 0, bm = 0;
                if (mh) { const int l = __ffs(mh) - 1; if (lane == l) bh = atomicAdd(hitCount, (uint32_t)__popc(mh)); bh = __shfl_sync(0xffffffffu, bh, l); }
                if (mm) { const int l = __ffs(mm) - 1; if (lane == l) bm = atomicAdd(missCount, (uint32_t)__popc(mm)); bm = __shfl_sync(0xffffffffu, bm, l); }
                if (finishedNow) {
                    const unsigned below = (1u << lane) - 1u;
                    if (tr.tri() >= 0) B.shadeQueue[bh + __popc(mh & below)] = pix;
                    else B.shadeQueue[B.capacity - 1u - (bm + __popc(mm & below))] = pix;
                }
            }
            const unsigned live = __ballot_sync(0xffffffffu, active);
            if (live == 0u) break;
            if (!exhausted && __popc(laneMask & ~live) >= refillIdle) break;
        }
    }
    if (COUNT && visitCounters) {
        for (int o = 16; o > 0; o >>= 1) { nodeVisits += __shfl_xor_sync(0xffffffffu, nodeVisits, o); leafTests += __shfl_xor_sync(0xffffffffu, leafTests, o); }
        if (lane == 0) { atomicAdd(visitCounters + 1, (unsigned long long)nodeVisits); atomicAdd(visitCounters + 2, (unsigned long long)leafTests); }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// k_pt_logic: apply the hit of the ray just traced, run the interactions of the segment(s), queue the next ray or
// write the pixel.  FIRST = the pass after the primary rays: one thread per pixel, the path is initialised here
// (same sampler values as k_pt_primary) and the segment-0 outputs (depth, motion vector, material mask) are written.
// Otherwise: one thread per entry of the shade queue (grid-stride over the device-side counts).
// ---------------------------------------------------------------------------------------------------------------------
VD void logic_body(const PathTraceParams& P, const WfBuffers& B, uint32_t pix, bool first, bool& emit, uint32_t& entry) {
    const int iy = (int)(pix / (uint32_t)P.width), ix = (int)(pix - (uint32_t)iy * (uint32_t)P.width);
    RayState r; int seg = 0; unsigned short mask = 0; Float2 sampleUv{0.0f, 0.0f};
    if (first) init_path(P, r, sampleUv, ix, iy);
    else load_state(B, pix, r, seg, mask);
    const float4 h = B.hit[pix];
    apply_hit(P, r, h.x, __float_as_int(h.w), h.y, h.z);
    if (first) {
        float outDepth = r.depth;
        mask = (unsigned short)r.matId;
        Float2 motion{0.0f, 0.0f};
        if (r.hit) {                                                            // HistoryCamera::WorldToScreenSpace src/kernel.cuh:144-151
            const Float3 v = mul(P.historyCamera.invCamMat, r.pos - P.historyCamera.pos);
            const Float2 ndc = (Float2{v.x, v.y} / v.z) / P.camera.tanHalfFov;
            motion = (Float2{0.5f, 0.5f} - ndc * Float2{0.5f, 0.5f}) - sampleUv;
        }
        motion = motion + 0.5f;
        if (outDepth != outDepth) outDepth = 0.0f;                              // NAN_DETECTER src/pathtrace.cuh:113-117
        if (motion.x != motion.x || motion.y != motion.y) motion = Float2{0.0f, 0.0f};
        P.depth[pix] = __float2half_rn(outDepth);
        P.motion[pix] = __float22half2_rn(make_float2(motion.x, motion.y));
    }
    emit = advance_path(P, r, seg, ix, iy, P.normal);
    if (emit) {
        r.isHitProcessed = false; save_state(B, pix, r, seg, mask);
        entry = pix | ((P.anyHitRays && (r.isShadowRay || seg == 4)) ? kAnyHitBit : 0u);
    } else finalize_path(P, r, mask, ix, iy);
}

__global__ void __launch_bounds__(kLogicThreads, RTRT_LOGIC_MINB) k_pt_logic_first(const PathTraceParams P, const WfBuffers B, uint32_t* countOut) {
    // 8x4-pixel warps like the primary kernel (sky / terrain coherence)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tilesX = (P.width + 7) / 8;
    const int tile = blockIdx.x * (kLogicThreads / 32) + warp;
    const int ix = (tile % tilesX) * 8 + (lane & 7), jc = (tile / tilesX) * 4 + (lane >> 3), jy = chunk_row(P, jc), iy = frame_row(P, jy);
    bool emit = false;
    const uint32_t pix = (uint32_t)((size_t)iy * P.width + ix);
    uint32_t entry = pix;
    if (ix < P.width && jc < P.chunkRows && jy < P.localRows) logic_body(P, B, pix, true, emit, entry);
    queue_append(emit, entry, B.rayQueue, countOut, lane);
}

__global__ void __launch_bounds__(kLogicThreads, RTRT_LOGIC_MINB) k_pt_logic(const PathTraceParams P, const WfBuffers B, const uint32_t* __restrict__ hitCount,
                                                            const uint32_t* __restrict__ missCount, uint32_t* countOut) {
    const uint32_t nHit = *hitCount, nMiss = *missCount;
    const uint32_t padHit = (nHit + 31u) & ~31u, total = padHit + nMiss;
    const int lane = threadIdx.x & 31;
    const uint32_t stride = gridDim.x * kLogicThreads;
    // every lane of a warp runs the same number of iterations so that queue_append stays warp-converged
    for (uint32_t q0 = blockIdx.x * kLogicThreads + (threadIdx.x & ~31u); q0 < total; q0 += stride) {
        const uint32_t q = q0 + lane;
        bool emit = false, valid;
        uint32_t slot;
        if (q0 < padHit) { valid = q < nHit; slot = q; }
        else { valid = q < total; slot = B.capacity - 1u - (q - padHit); }
        uint32_t pix = 0, entry = 0;
        if (valid) { pix = B.shadeQueue[slot]; logic_body(P, B, pix, false, emit, entry); }
        queue_append(emit, entry, B.rayQueue, countOut, lane);
    }
}

// rays traced = pixels + every queued ray (src/traverse.cuh:64-72: one TraverseBvh per passed guard)
__global__ void k_pt_count_rays(const uint32_t* __restrict__ counters, unsigned long long pixels, unsigned long long* out) {
    const unsigned long long n = pixels + counters[0] + counters[1] + counters[2] + counters[3];
    atomicAdd(out, n);                                                                  // one add per chunk of the frame
    atomicAdd(out + 4, n);                                                              // cumulative since rtrt_create (never reset)
}

SceneView make_scene_view(const Context& c);   // traverse.cu

static int ensure_wf_buffers(Context& c, WfBuffers& B) {
    const size_t px = (size_t)c.width * c.height;
    const size_t need = px * (6 * sizeof(float4) + 2 * sizeof(uint32_t)) + 256 * Context::kMaxChunks;
    if (c.wfBytes < need) {
        if (c.dWf) cudaFree(c.dWf);
        c.dWf = nullptr; c.wfBytes = 0;
        RTRT_CUDA(c, cudaMalloc(&c.dWf, need));
        c.wfBytes = need;
    }
    char* p = reinterpret_cast<char*>(c.dWf);
    B.counters = reinterpret_cast<uint32_t*>(p); p += 256 * Context::kMaxChunks;
    float4** f4[6] = {&B.st0, &B.st1, &B.st2, &B.st3, &B.st4, &B.hit};
    for (auto f : f4) { *f = reinterpret_cast<float4*>(p); p += px * sizeof(float4); }
    B.rayQueue = reinterpret_cast<uint32_t*>(p); p += px * sizeof(uint32_t);
    B.shadeQueue = reinterpret_cast<uint32_t*>(p);
    B.capacity = (uint32_t)px;
    return RTRT_OK;
}

void fill_pathtrace_params(Context& c, PathTraceParams& P);   // pathtrace.cu

// rows of chunk k when the context's local rows are dealt out in 8-row strips, strip s to chunk s % K
static int chunk_rows(int localRows, int K, int k) {
    if (K <= 1) return localRows;
    const int strips = (localRows + 7) / 8;
    const int mine = strips > k ? (strips - k + K - 1) / K : 0;
    int rows = mine * 8;
    if (mine > 0 && (strips - 1) % K == k) rows -= strips * 8 - localRows;
    return rows;
}

// One wavefront pipeline (primary, first logic, 4 x (trace, logic), ray count) over the pixels of chunk P.chunk, on `st`.
static void launch_chunk(Context& c, const PathTraceParams& P, const WfBuffers& B, cudaStream_t st, int smCount) {
    const int rows = P.chunkRows;
    const dim3 grid((c.width + kPtBlockX - 1) / kPtBlockX, (rows + kPtBlockY - 1) / kPtBlockY), block(kPtBlockX, kPtBlockY);
    const bool kt = c.kernelTiming && P.chunks == 1;
    if (kt) {
        for (auto& e : c.kev) if (!e) cudaEventCreate(&e);
        cudaEventRecord(c.kev[0], st);
    }
    const bool wide = c.wideBvh;
    if (wide && c.widePrimary) { if (c.countVisits) k_pt_primary<true, WideTrav><<<grid, block, 0, st>>>(P, B); else k_pt_primary<false, WideTrav><<<grid, block, 0, st>>>(P, B); }
    else      { if (c.countVisits) k_pt_primary<true, ParityTrav><<<grid, block, 0, st>>>(P, B); else k_pt_primary<false, ParityTrav><<<grid, block, 0, st>>>(P, B); }
    if (kt) cudaEventRecord(c.kev[1], st);
    const int tiles = ((c.width + 7) / 8) * ((rows + 3) / 4);
    k_pt_logic_first<<<(tiles + kLogicThreads / 32 - 1) / (kLogicThreads / 32), kLogicThreads, 0, st>>>(P, B, B.counters + 0);
    c.launches += 2;
    if (kt) cudaEventRecord(c.kev[2], st);
    // persistent grids sized for the chunk: at most one CTA per 128 pixels, never more than are resident at once
    if (!c.traceResident) {                          // per context = per device (a process may hold contexts on several GPUs)
        int perSm = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_pt_trace<false, ParityTrav>, kTraceThreads, 0);
        c.traceResident = smCount * (perSm > 0 ? perSm : 4);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_pt_trace<false, WideTrav>, kTraceThreads, 0);
        c.traceResidentWide = smCount * (perSm > 0 ? perSm : 4);
    }
    const int maxGrid = wide ? c.traceResidentWide : c.traceResident;
    const int want = (int)(((size_t)c.width * rows + 127) / 128);
    const int traceGrid = want < maxGrid ? (want > 0 ? want : 1) : maxGrid, logicGrid = traceGrid;
    if (wide && st == c.stream) wide_join(c);        // wide_bvh = 2: the collapse ran beside the primary-ray kernels; the bounce rounds need it
    for (int round = 0; round < 4; ++round) {
        uint32_t* cnt = B.counters;
        if (wide) {
            if (c.countVisits) k_pt_trace<true, WideTrav><<<traceGrid, kTraceThreads, 0, st>>>(P.sc, B, cnt + round, cnt + 8 + round, cnt + 16 + round, cnt + 24 + round, c.dRayCounter);
            else k_pt_trace<false, WideTrav><<<traceGrid, kTraceThreads, 0, st>>>(P.sc, B, cnt + round, cnt + 8 + round, cnt + 16 + round, cnt + 24 + round, nullptr);
        } else {
            if (c.countVisits) k_pt_trace<true, ParityTrav><<<traceGrid, kTraceThreads, 0, st>>>(P.sc, B, cnt + round, cnt + 8 + round, cnt + 16 + round, cnt + 24 + round, c.dRayCounter);
            else k_pt_trace<false, ParityTrav><<<traceGrid, kTraceThreads, 0, st>>>(P.sc, B, cnt + round, cnt + 8 + round, cnt + 16 + round, cnt + 24 + round, nullptr);
        }
        if (kt && round < 2) cudaEventRecord(c.kev[3 + 2 * round], st);
        k_pt_logic<<<logicGrid, kLogicThreads, 0, st>>>(P, B, cnt + 16 + round, cnt + 24 + round, cnt + round + 1);
        if (kt && round == 0) cudaEventRecord(c.kev[4], st);
        c.launches += 2;
    }
    k_pt_count_rays<<<1, 1, 0, st>>>(B.counters, (unsigned long long)c.width * rows, c.dRayCounter);
    c.launches += 1;
    if (kt) { cudaEventRecord(c.kev[6], st); c.kevPending = true; }
}

int launch_pathtrace_wavefront(Context& c) {
    PathTraceParams P{};
    fill_pathtrace_params(c, P);
    WfBuffers B{};
    int rc = ensure_wf_buffers(c, B);
    if (rc) return rc;
    const int smCount = c.smCount;
    RTRT_CUDA(c, cudaMemsetAsync(B.counters, 0, 256 * Context::kMaxChunks, c.stream));
    RTRT_CUDA(c, cudaMemsetAsync(c.dRayCounter, 0, 3 * sizeof(unsigned long long), c.stream));
    const int rows = local_rows(c);
    int K = c.ptChunks;
    if (K > (rows + 7) / 8) K = (rows + 7) / 8;
    if (K < 1) K = 1;
    if (c.wideBvh && (c.widePrimary || K > 1)) { int rcw = wide_join(c); if (rcw) return rcw; }
    if (K > 1 && !c.chunkFork) {
        RTRT_CUDA(c, cudaEventCreateWithFlags(&c.chunkFork, cudaEventDisableTiming));
        for (int k = 0; k < Context::kMaxChunks; ++k) {
            RTRT_CUDA(c, cudaStreamCreateWithFlags(&c.chunkStream[k], cudaStreamNonBlocking));
            RTRT_CUDA(c, cudaEventCreateWithFlags(&c.chunkJoin[k], cudaEventDisableTiming));
        }
    }
    if (K > 1) RTRT_CUDA(c, cudaEventRecord(c.chunkFork, c.stream));
    size_t queueOff = 0;
    for (int k = 0; k < K; ++k) {
        PathTraceParams Pk = P; WfBuffers Bk = B;
        Pk.chunks = K; Pk.chunk = k; Pk.chunkRows = chunk_rows(rows, K, k);
        const size_t chunkPx = (size_t)c.width * Pk.chunkRows;
        Bk.counters = B.counters + 64 * k;
        Bk.rayQueue = B.rayQueue + queueOff; Bk.shadeQueue = B.shadeQueue + queueOff; Bk.capacity = (uint32_t)chunkPx;
        queueOff += chunkPx;
        cudaStream_t st = c.stream;
        if (k > 0) { st = c.chunkStream[k]; RTRT_CUDA(c, cudaStreamWaitEvent(st, c.chunkFork, 0)); }
        launch_chunk(c, Pk, Bk, st, smCount);
        if (k > 0) { RTRT_CUDA(c, cudaEventRecord(c.chunkJoin[k], st)); RTRT_CUDA(c, cudaStreamWaitEvent(c.stream, c.chunkJoin[k], 0)); }
    }
    RTRT_CUDA(c, cudaGetLastError());
    return RTRT_OK;
}

}  // namespace rtrt
