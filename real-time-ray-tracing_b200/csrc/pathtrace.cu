// Stage 2 — unidirectional path tracer for sm_100a (rtrt_pathtrace).
//
// Semantics restated from src/pathtrace.cuh:11-128 (fixed 5-segment path: primary, <= 2 specular bounces, two
// diffuse/MIS interactions, environment lookup), src/raygen.cuh, src/traverse.cuh (RaySceneIntersect,
// UpdateMaterial), src/surfaceInteraction.cuh (Glossy/DiffuseSurfaceInteraction) and src/light.cuh:275-305.
// Same blue-noise indexing (sampleIdx = frameNum*4 + k, dims 0..3) so every pixel draws the reference's samples.
// Outputs are the reference's G-buffers (src/pathtrace.cuh:123-127, formats src/init.cu:473-500) as LINEAR
// device buffers: colour half3 + u16 material mask, normal half4, albedo half4, depth half, motion half2.
#include "pathtrace_common.cuh"

namespace rtrt {

template <bool COUNT>
__global__ void __launch_bounds__(kPtThreads) k_pathtrace(const PathTraceParams P) {
    __shared__ uint32_t sLink[16 * kPtThreads];
    __shared__ float sT[16 * kPtThreads];
    __shared__ unsigned int sRays;
    const int tid = threadIdx.y * kPtBlockX + threadIdx.x;
    if (tid == 0) sRays = 0;
    __syncthreads();
    const int ix = blockIdx.x * kPtBlockX + threadIdx.x;
    const int jy = blockIdx.y * kPtBlockY + threadIdx.y, iy = frame_row(P, jy);
    uint32_t rays = 0, nodeVisits = 0, leafTests = 0;
    if (ix < P.width && jy < P.localRows) {
        const float4 a = __ldg(P.sc.tlasRows), b = __ldg(P.sc.tlasRows + 1), c = __ldg(P.sc.tlasRows + 2);
        AABB sceneBox;
        sceneBox.max = Float3{sel_max(a.x, c.x), sel_max(a.y, c.y), sel_max(a.z, c.z)};
        sceneBox.min = Float3{sel_min(b.x, a.w), sel_min(b.y, b.w), sel_min(b.z, c.w)};
        uint32_t* stL = sLink + tid; float* stT = sT + tid;

        RayState r;
        r.beta0 = f3(1.0f); r.beta1 = f3(1.0f);
        r.hitLight = false; r.lightIdx = kDefaultLightId; r.isHitProcessed = true; r.isOccluded = false; r.isShadowRay = false;
        r.isDiffuse = false; r.hit = false; r.matId = 0; r.matType = 0; r.objectIdx = -1;
        r.normal = f3(0, -1, 0); r.fakeNormal = f3(0.0f); r.albedo = f3(1.0f); r.pos = f3(0.0f);
        r.rayConeWidth = 0.0f; r.depth = 0.0f; r.offset = 0.0f; r.normalDotRayDir = 0.0f; r.isRayIntoSurface = false;
        r.rayConeSpread = ray_cone_width(P.camera, ix, iy);
        Float4 rnd[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int si = P.frameNum * 4 + k;                                  // src/pathtrace.cuh:53-62
            rnd[k] = Float4{blue_rand(P.bn, ix, iy, si, 0), blue_rand(P.bn, ix, iy, si, 1), blue_rand(P.bn, ix, iy, si, 2), blue_rand(P.bn, ix, iy, si, 3)};
        }
        Float2 sampleUv;
        generate_ray(r.orig, r.dir, r.centerRaydir, sampleUv, P.camera, ix, iy, Float2{rnd[0].x, rnd[0].y}, Float2{rnd[0].z, rnd[0].w});
        // The reference unrolls five RaySceneIntersect calls (src/pathtrace.cuh:67-105).  Here they are ONE loop body so the
        // traversal and the diffuse interaction exist once in the instruction stream: the unrolled form is ~36 k SASS
        // instructions (580 KB) and spent 62 % of its issue slots waiting on instruction fetch (profiles/r1b).
        float outDepth = 0.0f; unsigned short mask = 0; Float2 motion{0.0f, 0.0f}; Float3 outNormal = f3(0.0f);
#pragma unroll 1
        for (int seg = 0; seg < 5; ++seg) {
            ray_scene_intersect<COUNT>(P, sceneBox, r, stL, stT, rays, nodeVisits, leafTests);
            if (seg == 0) {
                outDepth = r.depth;
                mask = (unsigned short)r.matId;
                if (r.hit) {                                                        // HistoryCamera::WorldToScreenSpace src/kernel.cuh:144-151
                    const Float3 v = mul(P.historyCamera.invCamMat, r.pos - P.historyCamera.pos);
                    const Float2 ndc = (Float2{v.x, v.y} / v.z) / P.camera.tanHalfFov;
                    motion = (Float2{0.5f, 0.5f} - ndc * Float2{0.5f, 0.5f}) - sampleUv;
                }
                motion = motion + 0.5f;
            }
            if (seg < 4) {
                const float g = seg == 0 ? rnd[0].x : (seg == 1 ? rnd[0].y : (seg == 2 ? rnd[0].z : rnd[0].w));
                glossy_interaction(r, g);
            }
            if (seg == 2 || seg == 3) {
                const bool first = seg == 2;
                Float3 beta = first ? r.beta1 : r.beta0;
                diffuse_interaction(first ? 0 : 1, P, r, beta, first ? rnd[0] : rnd[2], first ? rnd[1] : rnd[3]);
                if (first) { r.beta1 = beta; outNormal = r.fakeNormal; } else r.beta0 = beta;
            }
        }
        Float3 L0 = f3(0.0f);                                                   // GetLightSource src/light.cuh:275-305
        if (r.hitLight && !r.isOccluded && r.matType == MAT_SKY) L0 = env_light(P.env, r.dir);
        Float3 L2 = (L0 * r.beta0) * r.beta1;
        if (isnan3(L2)) L2 = f3(0.0f);                                          // NAN_DETECTER src/pathtrace.cuh:113-117
        if (isnan3(outNormal)) outNormal = f3(0.0f);
        if (outDepth != outDepth) outDepth = 0.0f;
        if (motion.x != motion.x || motion.y != motion.y) motion = Float2{0.0f, 0.0f};
        L2 = Float3{clampf(L2.x, 0.0f, 10.0f), clampf(L2.y, 0.0f, 10.0f), clampf(L2.z, 0.0f, 10.0f)};
        L2 = L2 / r.albedo;
        const size_t i = (size_t)iy * P.width + ix;
        P.color[i] = make_ushort4(f2h_bits(L2.x), f2h_bits(L2.y), f2h_bits(L2.z), mask);
        P.normal[i] = make_ushort4(f2h_bits(outNormal.x), f2h_bits(outNormal.y), f2h_bits(outNormal.z), f2h_bits(0.0f));
        P.albedo[i] = make_ushort4(f2h_bits(r.albedo.x), f2h_bits(r.albedo.y), f2h_bits(r.albedo.z), f2h_bits(0.0f));
        P.depth[i] = __float2half_rn(outDepth);
        P.motion[i] = __float22half2_rn(make_float2(motion.x, motion.y));
    }
    // rays traced: warp reduce -> one shared atomic per warp -> one global atomic per CTA
    for (int o = 16; o > 0; o >>= 1) rays += __shfl_xor_sync(0xffffffffu, rays, o);
    if ((tid & 31) == 0) atomicAdd(&sRays, rays);
    __syncthreads();
    if (tid == 0 && P.rayCounter) { atomicAdd(P.rayCounter, (unsigned long long)sRays); atomicAdd(P.rayCounter + 4, (unsigned long long)sRays); }
    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) { nodeVisits += __shfl_xor_sync(0xffffffffu, nodeVisits, o); leafTests += __shfl_xor_sync(0xffffffffu, leafTests, o); }
        if ((tid & 31) == 0 && P.rayCounter) { atomicAdd(P.rayCounter + 1, (unsigned long long)nodeVisits); atomicAdd(P.rayCounter + 2, (unsigned long long)leafTests); }
    }
}

SceneView make_scene_view(const Context& c);   // traverse.cu

void fill_pathtrace_params(Context& c, PathTraceParams& P) {
    P.sc = make_scene_view(c);
    P.materials = c.dMaterials; P.materialsIdx = c.dMaterialsIdx; P.numMaterials = c.numMaterials; P.numMaterialsIdx = c.numMaterialsIdx;
    P.camera = c.camera; P.historyCamera = c.historyCamera;
    P.env.sky = reinterpret_cast<const float4*>(c.dSky); P.env.skyCdf = c.dSkyCdf;
    P.env.sun = reinterpret_cast<const float4*>(c.dSun); P.env.sunCdf = c.dSunCdf;
    P.env.sunDir = c.sunDir; P.env.sunAngleCosThetaMax = c.sunAngleCosThetaMax;
    P.bn.sobol = c.dBlueNoise; P.bn.scramble = c.dBlueNoise + 65536; P.bn.rank = c.dBlueNoise + 65536 + 131072;
    for (int t = 0; t < 2; ++t) P.tex[t] = c.tex[t];
    P.frameNum = c.frameNum; P.width = c.width; P.height = c.height;
    P.rowBegin = c.rowBegin; P.rowEnd = c.rowEnd;
    P.stripH = c.stripH; P.stripShards = c.stripShards; P.stripShard = c.stripShard; P.localRows = local_rows(c);
    P.chunks = 1; P.chunk = 0; P.chunkRows = P.localRows;
    P.anyHitRays = c.anyHitRays ? 1 : 0;
    P.color = reinterpret_cast<ushort4*>(c.dColor); P.normal = reinterpret_cast<ushort4*>(c.dNormal);
    P.albedo = reinterpret_cast<ushort4*>(c.dAlbedo); P.depth = reinterpret_cast<__half*>(c.dDepth);
    P.motion = reinterpret_cast<__half2*>(c.dMotion);
    P.rayCounter = c.dRayCounter;
}

int launch_pathtrace_wavefront(Context& c);   // pathtrace_wf.cu

int local_rows(const Context& c) {
    if (c.stripShards <= 1) return c.rowEnd - c.rowBegin;
    const int strips = (c.height + c.stripH - 1) / c.stripH;                       // strips of the frame, the last one may be short
    const int mine = strips > c.stripShard ? (strips - c.stripShard + c.stripShards - 1) / c.stripShards : 0;
    int rows = mine * c.stripH;
    if (mine > 0 && (strips - 1) % c.stripShards == c.stripShard) rows -= strips * c.stripH - c.height;   // the short last strip is mine
    return rows;
}

int launch_pathtrace(Context& c) {
    if (!c.bvhValid) return c.fail(RTRT_ERR_STATE, "rtrt_pathtrace: BVH not built");
    if (!c.cameraSet) return c.fail(RTRT_ERR_STATE, "rtrt_pathtrace: camera not set");
    // the path-trace kernels keep the reference's 16-entry stack (src/traverse.h:28) in shared memory; a deeper one exists only in
    // rtrt_traverse (<= 64) and in the wide layout (option wide_bvh: spill stack, no cap) — say so instead of clamping silently
    if (!c.wideBvh && c.stackSize > 16)
        return c.fail(RTRT_ERR_ARG, "rtrt_pathtrace: stack_size > 16 is honoured by rtrt_traverse only (use option wide_bvh for deep scenes)");
    int rc = ensure_frame_resources(c);       // G-buffers, blue noise, default materials / textures, sky
    if (rc) return rc;
    rc = ensure_wide(c);
    if (rc) return rc;
    if (c.wavefront) return launch_pathtrace_wavefront(c);
    if (c.wideBvh) return c.fail(RTRT_ERR_ARG, "rtrt_pathtrace: the megakernel (wavefront = 0) traverses the reference layout only; switch wide_bvh off");
    PathTraceParams P{};
    fill_pathtrace_params(c, P);
    RTRT_CUDA(c, cudaMemsetAsync(c.dRayCounter, 0, 3 * sizeof(unsigned long long), c.stream));
    const int rows = local_rows(c);
    const dim3 grid((c.width + kPtBlockX - 1) / kPtBlockX, (rows + kPtBlockY - 1) / kPtBlockY), block(kPtBlockX, kPtBlockY);
    if (c.countVisits) k_pathtrace<true><<<grid, block, 0, c.stream>>>(P);
    else k_pathtrace<false><<<grid, block, 0, c.stream>>>(P);
    c.launches += 1;
    RTRT_CUDA(c, cudaGetLastError());
    return RTRT_OK;
}

}  // namespace rtrt
