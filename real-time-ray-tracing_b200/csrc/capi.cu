// C ABI implementation (include/rtrt.h): context lifetime, mesh upload, stage entry points and the
// observation-point getters.  Host-side mirror of RayTracer::init / draw / cleanup
// (src/init.cu:53-439, src/kernel.cu:259-398, src/init.cu:601-663) for the hot path only.
#include "context.h"
#include <cstring>
#include <cstdlib>
#include <cstdio>
#include <vector>
#include <new>

using namespace rtrt;


namespace rtrt {

template <typename T>
static cudaError_t dalloc(T*& p, size_t count) {
    p = nullptr;
    return cudaMalloc((void**)&p, count ? count * sizeof(T) : sizeof(T));
}
template <typename T>
static void dfree(T*& p) { if (p) cudaFree(p); p = nullptr; }

static void free_mesh(Context& c) {
    if (c.wideStream) cudaStreamSynchronize(c.wideStream);
    if (c.wideTlasStream) cudaStreamSynchronize(c.wideTlasStream);
    c.widePending = false;
    if (c.ownsMesh) { dfree(c.dPos); dfree(c.dNrm); dfree(c.dIdx); }
    c.dPos = nullptr; c.dNrm = nullptr; c.dIdx = nullptr; c.ownsMesh = false;
    dfree(c.dTriangles); dfree(c.dAabbs); dfree(c.dMorton); dfree(c.dReorderIdx); dfree(c.dNodes); dfree(c.dMortonPre);
    dfree(c.dTlasAabbs); dfree(c.dTlasMorton); dfree(c.dTlasReorderIdx); dfree(c.dTlasNodes); dfree(c.dTlasMortonPre);
    if (c.dSortTmp) { cudaFree(c.dSortTmp); c.dSortTmp = nullptr; c.sortTmpBytes = 0; }
    if (c.dPresort) { cudaFree(c.dPresort); c.dPresort = nullptr; c.presortBytes = 0; }
    dfree(c.dIdxSorted); dfree(c.dTriPerm);
    if (c.dWideNodes) cudaFree(c.dWideNodes); if (c.dWideTris) cudaFree(c.dWideTris); if (c.dWideQueue) cudaFree(c.dWideQueue);
    c.dWideNodes = nullptr; c.dWideTris = nullptr; c.dWideQueue = nullptr; c.wideNodeCap = 0; c.wideValid = false; c.wideStatPending = false;
    c.numVerts = c.numTris = c.numBatches = c.numBatchesPadded = 0;
    c.bvhValid = false;
}

// Allocation block of RayTracer::init (src/init.cu:159-187).  Every per-triangle array holds B*1024
// entries (the reference allocates ceil4(N) for some of them; rounding up to whole batches lets every
// CTA write full rows) and is zero-filled once so untouched tails are deterministic.
static int alloc_stage1(Context& c, uint32_t V, uint32_t N) {
    c.numVerts = V; c.numTris = N;
    c.numBatches = (N + kBatchSize - 1) / kBatchSize;                        // src/init.cu:117-126
    c.numBatchesPadded = (c.numBatches + kBatchSize - 1) / kBatchSize * kBatchSize;
    const size_t n = (size_t)c.numBatches * kBatchSize, bp = c.numBatchesPadded;
    RTRT_CUDA(c, dalloc(c.dTriangles, n)); RTRT_CUDA(c, dalloc(c.dAabbs, n));
    RTRT_CUDA(c, dalloc(c.dMorton, n));    RTRT_CUDA(c, dalloc(c.dReorderIdx, n));
    RTRT_CUDA(c, dalloc(c.dNodes, n));     RTRT_CUDA(c, dalloc(c.dMortonPre, n));
    RTRT_CUDA(c, dalloc(c.dTlasAabbs, bp)); RTRT_CUDA(c, dalloc(c.dTlasMorton, bp)); RTRT_CUDA(c, dalloc(c.dTlasReorderIdx, bp));
    RTRT_CUDA(c, dalloc(c.dTlasNodes, bp)); RTRT_CUDA(c, dalloc(c.dTlasMortonPre, bp));
    RTRT_CUDA(c, cudaMemsetAsync(c.dTriangles, 0, n * sizeof(Triangle), c.stream));
    RTRT_CUDA(c, cudaMemsetAsync(c.dAabbs, 0, n * sizeof(AABB), c.stream));
    RTRT_CUDA(c, cudaMemsetAsync(c.dNodes, 0, n * sizeof(BVHNode), c.stream));
    RTRT_CUDA(c, cudaMemsetAsync(c.dMorton, 0xFF, n * sizeof(uint32_t), c.stream));       // src/init.cu:167
    RTRT_CUDA(c, cudaMemsetAsync(c.dMortonPre, 0xFF, n * sizeof(uint32_t), c.stream));
    RTRT_CUDA(c, cudaMemsetAsync(c.dReorderIdx, 0, n * sizeof(uint32_t), c.stream));
    RTRT_CUDA(c, cudaMemsetAsync(c.dTlasAabbs, 0, bp * sizeof(AABB), c.stream));
    RTRT_CUDA(c, cudaMemsetAsync(c.dTlasNodes, 0, bp * sizeof(BVHNode), c.stream));
    RTRT_CUDA(c, cudaMemsetAsync(c.dTlasMorton, 0xFF, bp * sizeof(uint32_t), c.stream));  // src/init.cu:178
    RTRT_CUDA(c, cudaMemsetAsync(c.dTlasMortonPre, 0xFF, bp * sizeof(uint32_t), c.stream));
    RTRT_CUDA(c, cudaMemsetAsync(c.dTlasReorderIdx, 0, bp * sizeof(uint32_t), c.stream));
    return RTRT_OK;
}

// GenerateSmoothNormals (src/kernel.cu:228-257): angle-weighted area normals accumulated with float atomics.
__device__ __forceinline__ float dev_dot(Float3 a, Float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ Float3 dev_sub(Float3 a, Float3 b) { return Float3{a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ float angle_between(Float3 a, Float3 b) {           // src/linearMath.h:497
    return acosf(dev_dot(a, b) / sqrtf(dev_dot(a, a) * dev_dot(b, b)));
}
__global__ void k_smooth_normals(const uint32_t* __restrict__ idx, const float* __restrict__ pos, float* nrm, uint32_t N) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const uint32_t i0 = idx[3 * i], i1 = idx[3 * i + 1], i2 = idx[3 * i + 2];
    const Float3 v0{pos[3 * i0], pos[3 * i0 + 1], pos[3 * i0 + 2]}, v1{pos[3 * i1], pos[3 * i1 + 1], pos[3 * i1 + 2]},
        v2{pos[3 * i2], pos[3 * i2 + 1], pos[3 * i2 + 2]};
    const Float3 a = dev_sub(v2, v0), b = dev_sub(v2, v1);
    const Float3 n{(a.y * b.z - a.z * b.y) / 2, (a.z * b.x - a.x * b.z) / 2, (a.x * b.y - a.y * b.x) / 2};
    const float w0 = angle_between(dev_sub(v2, v0), dev_sub(v1, v0));
    const float w1 = angle_between(dev_sub(v2, v1), dev_sub(v0, v1));
    const float w2 = angle_between(dev_sub(v0, v2), dev_sub(v1, v2));
    atomicAdd(nrm + 3 * i0, n.x * w0); atomicAdd(nrm + 3 * i0 + 1, n.y * w0); atomicAdd(nrm + 3 * i0 + 2, n.z * w0);
    atomicAdd(nrm + 3 * i1, n.x * w1); atomicAdd(nrm + 3 * i1 + 1, n.y * w1); atomicAdd(nrm + 3 * i1 + 2, n.z * w1);
    atomicAdd(nrm + 3 * i2, n.x * w2); atomicAdd(nrm + 3 * i2 + 1, n.y * w2); atomicAdd(nrm + 3 * i2 + 2, n.z * w2);
}

static int ensure_ray_stage(Context& c, uint32_t n) {
    if (n <= c.rayStageCap) return RTRT_OK;
    dfree(c.dRayStage); dfree(c.dHitStage); c.rayStageCap = 0;
    RTRT_CUDA(c, dalloc(c.dRayStage, n)); RTRT_CUDA(c, dalloc(c.dHitStage, n));
    c.rayStageCap = n;
    return RTRT_OK;
}

}  // namespace rtrt

#define CTX_OR_FAIL(ctx) do { if (!(ctx)) return RTRT_ERR_ARG; } while (0)

static void rtrt_set_sky_params_defaults(rtrt_sky_params* p) {       // src/settingParams.h:41-45
    p->timeOfDay = 0.25f; p->sunAxisAngle = 45.0f; p->skyScalar = 0.01f; p->sunScalar = 0.01f; p->sunAngle = 0.6f;
}
static void rtrt_postprocess_defaults(rtrt_postprocess_params* p) {          // src/settingParams.h:92-99, 112-116
    p->exposure = 1.0f; p->gain = 40.0f; p->maxWhite = 7.0f; p->gamma = 2.2f;
    p->enablePostProcess = 1; p->enableAutoExposure = 1; p->enableSharpening = 1; p->enableToneMapping = 1;
}

static void rtrt_set_denoise_defaults(rtrt_denoise_params* d, rtrt_pass_settings* s) {   // src/settingParams.h:144-157, 87-91
    d->local_denoise_sigma_normal = 100.0f; d->local_denoise_sigma_depth = 0.1f; d->local_denoise_sigma_material = 100.0f;
    d->large_denoise_sigma_normal = 100.0f; d->large_denoise_sigma_depth = 0.01f; d->large_denoise_sigma_material = 100.0f;
    d->temporal_denoise_sigma_normal = 100.0f; d->temporal_denoise_sigma_depth = 0.1f; d->temporal_denoise_sigma_material = 100.0f;
    d->noise_threshold_local = 0.001f; d->noise_threshold_large = 0.001f;
    s->enableTemporalDenoising = 1; s->enableLocalSpatialFilter = 1; s->enableWideSpatialFilter = 1; s->enableTemporalDenoising2 = 1;
}

static HistoryCamera history_setup(const Camera& cam) {                // HistoryCamera::Setup src/kernel.cuh:137-142
    HistoryCamera h;
    h.invCamMat = Mat3{cam.left.x, cam.up.x, cam.dir.x, cam.left.y, cam.up.y, cam.dir.y, cam.left.z, cam.up.z, cam.dir.z};
    h.pos = cam.pos;
    return h;
}

extern "C" {

int rtrt_abi_version(void) { return RTRT_ABI_VERSION; }

int rtrt_create(const rtrt_config* cfg, int device, rtrt_ctx** out) {
    if (!out) return RTRT_ERR_ARG;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) return RTRT_ERR_CUDA;   // no CPU fallback: fail loudly
    if (device < 0 || device >= ndev) return RTRT_ERR_ARG;
    rtrt_ctx* h = new (std::nothrow) rtrt_ctx();
    if (!h) return RTRT_ERR_CAPACITY;
    Context& c = h->c;
    c.device = device;
    if (cfg) c.cfg = *cfg; else rtrt_config_defaults(&c.cfg);
    c.width = c.cfg.width; c.height = c.cfg.height;
    rtrt_set_sky_params_defaults(&c.skyParams);
    rtrt_set_denoise_defaults(&c.denoiseParams, &c.passSettings);
    rtrt_postprocess_defaults(&c.postParams);
    if (cudaSetDevice(device) != cudaSuccess) { delete h; return RTRT_ERR_CUDA; }
    if (cudaDeviceGetAttribute(&c.smCount, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || c.smCount <= 0) c.smCount = 148;
    if (cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking) != cudaSuccess) { delete h; return RTRT_ERR_CUDA; }
    if (cudaStreamCreateWithFlags(&c.copyStream, cudaStreamNonBlocking) != cudaSuccess) { delete h; return RTRT_ERR_CUDA; }
    if (cudaStreamCreateWithFlags(&c.uploadStream, cudaStreamNonBlocking) != cudaSuccess) { delete h; return RTRT_ERR_CUDA; }
    for (cudaEvent_t* e : {&c.evGeomReady, &c.evGeomFree, &c.evStaged, &c.evReadDone}) cudaEventCreateWithFlags(e, cudaEventDisableTiming);
    for (auto& e : c.ev) { cudaEventCreate(&e[0]); cudaEventCreate(&e[1]); }
    if (cudaMalloc((void**)&c.dTicket, 64) != cudaSuccess || cudaMemset(c.dTicket, 0, 64) != cudaSuccess) { delete h; return RTRT_ERR_CUDA; }
    *out = h;
    return RTRT_OK;
}

int rtrt_destroy(rtrt_ctx* ctx) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    cudaSetDevice(c.device);
    cudaStreamSynchronize(c.stream);
    if (c.copyStream) cudaStreamSynchronize(c.copyStream);
    if (c.uploadStream) cudaStreamSynchronize(c.uploadStream);
    multi_release(c);
    free_mesh(c);
    free_frame_resources(c);
    dfree(c.dStage);
    dfree(c.dRayStage); dfree(c.dHitStage); dfree(c.dTicket);
    for (auto& e : c.ev) { if (e[0]) cudaEventDestroy(e[0]); if (e[1]) cudaEventDestroy(e[1]); }
    for (auto& e : c.kev) if (e) cudaEventDestroy(e);
    if (c.evWideBegin) cudaEventDestroy(c.evWideBegin);
    for (auto& st : c.chunkStream) if (st) cudaStreamDestroy(st);
    for (auto& ev : c.chunkJoin) if (ev) cudaEventDestroy(ev);
    if (c.chunkFork) cudaEventDestroy(c.chunkFork);
    if (c.evWideFork) cudaEventDestroy(c.evWideFork);
    if (c.evWideDone) cudaEventDestroy(c.evWideDone);
    if (c.wideStream) cudaStreamDestroy(c.wideStream);
    if (c.evTlasFork) cudaEventDestroy(c.evTlasFork);
    if (c.evTlasDone) cudaEventDestroy(c.evTlasDone);
    if (c.wideTlasStream) cudaStreamDestroy(c.wideTlasStream);
    for (cudaEvent_t e : {c.evGeomReady, c.evGeomFree, c.evStaged, c.evReadDone}) if (e) cudaEventDestroy(e);
    if (c.copyStream) cudaStreamDestroy(c.copyStream);
    if (c.uploadStream) cudaStreamDestroy(c.uploadStream);
    if (c.stream) cudaStreamDestroy(c.stream);
    delete ctx;
    return RTRT_OK;
}

const char* rtrt_last_error(const rtrt_ctx* ctx) { return ctx ? ctx->c.lastError.c_str() : "null context"; }

int rtrt_set_mesh(rtrt_ctx* ctx, const float* pos, const float* nrm, const uint32_t* idx, uint32_t V, uint32_t N) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (!pos || !idx || V == 0 || N == 0) return c.fail(RTRT_ERR_ARG, "rtrt_set_mesh: null/empty input");
    if ((uint64_t)N > (1ull << 30)) return c.fail(RTRT_ERR_CAPACITY, "rtrt_set_mesh: more than 2^30 triangles");
    for (uint32_t i = 0; i < 3u * N; ++i)
        if (idx[i] >= V) return c.fail(RTRT_ERR_ARG, "rtrt_set_mesh: index out of range");
    RTRT_CUDA(c, cudaSetDevice(c.device));
    RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    free_mesh(c);
    c.ownsMesh = true;
    RTRT_CUDA(c, dalloc(c.dPos, (size_t)3 * V)); RTRT_CUDA(c, dalloc(c.dNrm, (size_t)3 * V)); RTRT_CUDA(c, dalloc(c.dIdx, (size_t)3 * N));
    RTRT_CUDA(c, cudaMemcpyAsync(c.dPos, pos, (size_t)3 * V * sizeof(float), cudaMemcpyHostToDevice, c.stream));
    RTRT_CUDA(c, cudaMemcpyAsync(c.dIdx, idx, (size_t)3 * N * sizeof(uint32_t), cudaMemcpyHostToDevice, c.stream));
    if (nrm) {
        RTRT_CUDA(c, cudaMemcpyAsync(c.dNrm, nrm, (size_t)3 * V * sizeof(float), cudaMemcpyHostToDevice, c.stream));
    } else {
        RTRT_CUDA(c, cudaMemsetAsync(c.dNrm, 0, (size_t)3 * V * sizeof(float), c.stream));
        k_smooth_normals<<<(N + 63) / 64, 64, 0, c.stream>>>(c.dIdx, c.dPos, c.dNrm, N);   // src/kernel.cu:315
        c.launches += 1;
    }
    int rc = alloc_stage1(c, V, N);
    if (rc) return rc;
    RTRT_CUDA(c, cudaStreamSynchronize(c.stream));     // host arrays are never retained
    return RTRT_OK;
}

int rtrt_set_mesh_device(rtrt_ctx* ctx, const float* d_pos, const float* d_nrm, const uint32_t* d_idx, uint32_t V, uint32_t N) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (!d_pos || !d_nrm || !d_idx || V == 0 || N == 0) return c.fail(RTRT_ERR_ARG, "rtrt_set_mesh_device: null/empty input");
    RTRT_CUDA(c, cudaSetDevice(c.device));
    RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    free_mesh(c);
    c.ownsMesh = false;
    c.dPos = const_cast<float*>(d_pos); c.dNrm = const_cast<float*>(d_nrm); c.dIdx = const_cast<uint32_t*>(d_idx);
    int rc = alloc_stage1(c, V, N);
    if (rc) return rc;
    RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    return RTRT_OK;
}

int rtrt_update_vertices(rtrt_ctx* ctx, const float* pos, const float* nrm, uint32_t V) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (!pos) return c.fail(RTRT_ERR_ARG, "rtrt_update_vertices: null positions");
    if (!c.dPos || c.numTris == 0) return c.fail(RTRT_ERR_STATE, "rtrt_update_vertices: no mesh set");
    if (!c.ownsMesh) return c.fail(RTRT_ERR_STATE, "rtrt_update_vertices: mesh buffers are caller-owned (rtrt_set_mesh_device)");
    if (V != c.numVerts) return c.fail(RTRT_ERR_ARG, "rtrt_update_vertices: vertex count differs from the mesh");
    RTRT_CUDA(c, cudaSetDevice(c.device));
    // upload stream: starts once the last build has consumed the previous contents, overlaps whatever the main stream runs
    // (path trace / denoise of the frame in flight read the Triangle array, not the vertex buffers)
    if (c.geomFreeValid) RTRT_CUDA(c, cudaStreamWaitEvent(c.uploadStream, c.evGeomFree, 0));
    RTRT_CUDA(c, cudaMemcpyAsync(c.dPos, pos, (size_t)3 * V * sizeof(float), cudaMemcpyHostToDevice, c.uploadStream));
    if (nrm) RTRT_CUDA(c, cudaMemcpyAsync(c.dNrm, nrm, (size_t)3 * V * sizeof(float), cudaMemcpyHostToDevice, c.uploadStream));
    RTRT_CUDA(c, cudaEventRecord(c.evGeomReady, c.uploadStream));
    c.geomPending = true;
    c.bvhValid = false;
    return RTRT_OK;
}

}  // extern "C" (helpers below have C++ linkage)
namespace rtrt {
int wait_geometry(Context& c) {
    if (c.geomPending) { RTRT_CUDA(c, cudaStreamWaitEvent(c.stream, c.evGeomReady, 0)); c.geomPending = false; }
    return RTRT_OK;
}
// After a stream synchronisation: did the top-level CTA of a fused build give up waiting for the batch boxes (bvh_build.cu)?
// The BVH is then incomplete: mark it invalid and report instead of tracing against a stale top level.
int check_build_error(Context& c) {
    if (!c.buildErrPending) return RTRT_OK;
    c.buildErrPending = false;
    uint32_t w[2] = {0, 0};
    RTRT_CUDA(c, cudaMemcpy(w, c.dTicket, sizeof(w), cudaMemcpyDeviceToHost));
    if (w[1] == 0u) return RTRT_OK;
    RTRT_CUDA(c, cudaMemset(c.dTicket + 1, 0, sizeof(uint32_t)));
    c.bvhValid = false;
    return c.fail(RTRT_ERR_CUDA, "rtrt_build_bvh: the top-level CTA timed out waiting for the batch boxes; the BVH is incomplete (rebuild)");
}
int geometry_consumed(Context& c) {
    RTRT_CUDA(c, cudaEventRecord(c.evGeomFree, c.stream));
    c.geomFreeValid = true;
    return RTRT_OK;
}
}  // namespace rtrt
extern "C" {

int rtrt_set_option_i(rtrt_ctx* ctx, const char* name, int64_t value) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (!name) return c.fail(RTRT_ERR_ARG, "rtrt_set_option_i: null name");
    if (!strcmp(name, "keep_presort_morton")) { c.keepPresortMorton = value != 0; return RTRT_OK; }
    if (!strcmp(name, "stack_size")) { if (value < 1 || value > 64) return c.fail(RTRT_ERR_ARG, "stack_size must be 1..64"); c.stackSize = (int)value; return RTRT_OK; }
    if (!strcmp(name, "assume_bvh_valid")) { c.bvhValid = value != 0; c.wideValid = false; return RTRT_OK; }
    if (!strcmp(name, "wide_bvh")) {
        if (value < 0 || value > 2) return c.fail(RTRT_ERR_ARG, "wide_bvh must be 0, 1 or 2");
        { int rc = wide_join(c); if (rc) return rc; }
        c.wideBvh = value != 0; c.widePrimary = value != 2; return RTRT_OK;
    }
    if (!strcmp(name, "morton_bits")) { if (value != 60 && value != 63) return c.fail(RTRT_ERR_ARG, "morton_bits must be 60 or 63"); c.mortonBits = (int)value; c.bvhValid = false; return RTRT_OK; }
    if (!strcmp(name, "global_sort")) { c.globalSort = value != 0; c.bvhValid = false; return RTRT_OK; }
    if (!strcmp(name, "wavefront")) { c.wavefront = value != 0; return RTRT_OK; }
    if (!strcmp(name, "count_visits")) { c.countVisits = value != 0; return RTRT_OK; }
    if (!strcmp(name, "any_hit_rays")) { c.anyHitRays = value != 0; return RTRT_OK; }
    if (!strcmp(name, "kernel_timing")) { c.kernelTiming = value != 0; return RTRT_OK; }
    if (!strcmp(name, "pt_chunks")) { if (value < 1 || value > Context::kMaxChunks) return c.fail(RTRT_ERR_ARG, "pt_chunks must be 1..8"); c.ptChunks = (int)value; return RTRT_OK; }
    if (!strcmp(name, "tlas_spin_limit")) { if (value < 0 || value > (1ll << 31)) return c.fail(RTRT_ERR_ARG, "tlas_spin_limit out of range"); c.tlasSpinLimit = (uint32_t)value; return RTRT_OK; }
    if (!strcmp(name, "keep_all_triangles")) { c.keepAllTris = value != 0; c.bvhValid = false; return RTRT_OK; }
    if (!strcmp(name, "max_loop")) { if (value < 1) return c.fail(RTRT_ERR_ARG, "max_loop must be >= 1"); c.maxLoop = (int)value; return RTRT_OK; }
    return c.fail(RTRT_ERR_ARG, "rtrt_set_option_i: unknown option");
}

int rtrt_build_bvh(rtrt_ctx* ctx) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    RTRT_CUDA(c, cudaSetDevice(c.device));
    { int rc = wide_join(c); if (rc) return rc; }            // a collapse still running on the side stream reads the arrays this build rewrites
    RTRT_CUDA(c, cudaEventRecord(c.ev[STAGE_BUILD][0], c.stream));
    int rc = launch_build_bvh(c);
    if (rc) return rc;
    c.wideValid = false;
    c.evWidePending = false;
    if (c.wideBvh) {                                                      // part of rtrt_build_bvh when the option is on; timed separately (wide_build_ms)
        if (!c.evWideBegin) RTRT_CUDA(c, cudaEventCreate(&c.evWideBegin));
        RTRT_CUDA(c, cudaEventRecord(c.evWideBegin, c.stream));
        c.evWidePending = c.widePrimary;                                  // wide_bvh = 2: the collapse is on the side stream, outside this stage
        rc = launch_wide_build(c); if (rc) return rc;
    }
    RTRT_CUDA(c, cudaEventRecord(c.ev[STAGE_BUILD][1], c.stream));
    c.bvhValid = true;
    c.evPending[STAGE_BUILD] = true;    // resolved lazily in rtrt_get_timings
    return RTRT_OK;
}

int rtrt_traverse(rtrt_ctx* ctx, const rtrt_ray* d_rays, uint32_t n, rtrt_hit* d_hits, int any_hit) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if ((!d_rays || !d_hits) && n) return c.fail(RTRT_ERR_ARG, "rtrt_traverse: null buffer");
    RTRT_CUDA(c, cudaSetDevice(c.device));
    c.timings.rays_traced = n; c.raysOnDevice = false;
    RTRT_CUDA(c, cudaEventRecord(c.ev[STAGE_TRAVERSE][0], c.stream));
    int rc = launch_traverse(c, reinterpret_cast<const Ray*>(d_rays), n, reinterpret_cast<Hit*>(d_hits), any_hit, nullptr);
    if (rc) return rc;
    RTRT_CUDA(c, cudaEventRecord(c.ev[STAGE_TRAVERSE][1], c.stream));
    c.evPending[STAGE_TRAVERSE] = true;
    return RTRT_OK;
}

int rtrt_traverse_host(rtrt_ctx* ctx, const rtrt_ray* h_rays, uint32_t n, rtrt_hit* h_hits, int any_hit) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if ((!h_rays || !h_hits) && n) return c.fail(RTRT_ERR_ARG, "rtrt_traverse_host: null buffer");
    if (n == 0) return RTRT_OK;
    RTRT_CUDA(c, cudaSetDevice(c.device));
    int rc = ensure_ray_stage(c, n);
    if (rc) return rc;
    RTRT_CUDA(c, cudaMemcpyAsync(c.dRayStage, h_rays, (size_t)n * sizeof(Ray), cudaMemcpyHostToDevice, c.stream));
    rc = launch_traverse(c, c.dRayStage, n, c.dHitStage, any_hit, nullptr);
    if (rc) return rc;
    RTRT_CUDA(c, cudaMemcpyAsync(h_hits, c.dHitStage, (size_t)n * sizeof(Hit), cudaMemcpyDeviceToHost, c.stream));
    RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    c.timings.rays_traced = n; c.raysOnDevice = false;
    return RTRT_OK;
}

int rtrt_get_buffer(rtrt_ctx* ctx, int which, void** dptr, size_t* bytes) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (!dptr || !bytes) return c.fail(RTRT_ERR_ARG, "rtrt_get_buffer: null out pointer");
    if (which >= RTRT_BUF_SKY && which <= RTRT_BUF_SUN_CDF && (c.skyDirty || !c.dSky)) {   // lazy regeneration, src/kernel.cu:289
        int rc = ensure_frame_resources(c);
        if (rc) return rc;
    }
    if (which >= RTRT_BUF_WIDE_NODES && which <= RTRT_BUF_WIDE_STAT) { int rc = wide_join(c); if (rc) return rc; }
    const size_t n = (size_t)c.numBatches * kBatchSize, bp = c.numBatchesPadded, px = (size_t)c.gbufW * c.gbufH;
    void* p = nullptr; size_t b = 0;
    switch (which) {
        case RTRT_BUF_TRIANGLES: p = c.dTriangles; b = n * sizeof(Triangle); break;
        case RTRT_BUF_AABBS: p = c.dAabbs; b = n * sizeof(AABB); break;
        case RTRT_BUF_MORTON_PRESORT: p = c.dMortonPre; b = n * 4; break;
        case RTRT_BUF_MORTON: p = c.dMorton; b = n * 4; break;
        case RTRT_BUF_REORDER_IDX: p = c.dReorderIdx; b = n * 4; break;
        case RTRT_BUF_BVH_NODES: p = c.dNodes; b = n * sizeof(BVHNode); break;
        case RTRT_BUF_TLAS_AABBS: p = c.dTlasAabbs; b = bp * sizeof(AABB); break;
        case RTRT_BUF_TLAS_MORTON_PRESORT: p = c.dTlasMortonPre; b = bp * 4; break;
        case RTRT_BUF_TLAS_MORTON: p = c.dTlasMorton; b = bp * 4; break;
        case RTRT_BUF_TLAS_REORDER_IDX: p = c.dTlasReorderIdx; b = bp * 4; break;
        case RTRT_BUF_TLAS_BVH_NODES: p = c.dTlasNodes; b = bp * sizeof(BVHNode); break;
        case RTRT_BUF_TRI_PERMUTATION: p = c.dTriPerm; b = (size_t)c.numTris * 4; break;
        case RTRT_BUF_RENDER_COLOR: p = c.dColor; b = px * 8; break;
        case RTRT_BUF_NORMAL: p = c.dNormal; b = px * 8; break;
        case RTRT_BUF_ALBEDO: p = c.dAlbedo; b = px * 8; break;
        case RTRT_BUF_DEPTH: p = c.dDepth; b = px * 2; break;
        case RTRT_BUF_MOTION_VECTOR: p = c.dMotion; b = px * 4; break;
        case RTRT_BUF_ACCUMULATION_COLOR: p = c.dAccum; b = px * 8; break;
        case RTRT_BUF_HISTORY_COLOR: p = c.dHistoryColor; b = px * 8; break;
        case RTRT_BUF_HISTORY_DEPTH: p = c.dHistoryDepth; b = px * 2; break;
        case RTRT_BUF_NOISE_LEVEL: p = c.dNoise; b = (size_t)((c.gbufW + 7) / 8) * ((c.gbufH + 7) / 8) * 2; break;
        case RTRT_BUF_NOISE_LEVEL_16: p = c.dNoise16; b = (size_t)((c.gbufW + 15) / 16) * ((c.gbufH + 15) / 16) * 2; break;
        case RTRT_BUF_ACCUM_RADIANCE: p = c.dAccumRadiance; b = px * 16; break;
        case RTRT_BUF_OUTPUT_RGBA8: p = c.dOutRgba; b = (size_t)c.ppOutW * c.ppOutH * 4; break;
        case RTRT_BUF_COLOR64: p = c.dColor64; b = (size_t)c.ppW64 * c.ppH64 * 8; break;
        case RTRT_BUF_EXPOSURE: p = c.dExposure; b = c.dExposure ? 16 : 0; break;
        case RTRT_BUF_HISTOGRAM: p = c.dExposure ? (void*)((float*)c.dExposure + 4) : nullptr; b = c.dExposure ? 256 : 0; break;
        case RTRT_BUF_WF_COUNTERS: p = c.dWf; b = c.dWf ? 256 : 0; break;
        case RTRT_BUF_WIDE_NODES: p = c.wideValid ? c.dWideNodes : nullptr; b = ((size_t)c.numBatches * 512 + 2 * (size_t)c.numBatches + 8) * 128; break;
        case RTRT_BUF_WIDE_TRIS: p = c.wideValid ? c.dWideTris : nullptr; b = n * 48; break;
        case RTRT_BUF_WIDE_STAT: p = c.wideValid ? (void*)((char*)c.dWideQueue + (2 * (size_t)c.numBatches + 8) * 40) : nullptr; b = 16; break;
        case RTRT_BUF_SKY: p = c.dSky; b = sizeof(float) * 4 * kSkySize; break;
        case RTRT_BUF_SKY_CDF: p = c.dSkyCdf; b = sizeof(float) * kSkySize; break;
        case RTRT_BUF_SUN: p = c.dSun; b = sizeof(float) * 4 * kSunSize; break;
        case RTRT_BUF_SUN_CDF: p = c.dSunCdf; b = sizeof(float) * kSunSize; break;
        default: return c.fail(RTRT_ERR_ARG, "rtrt_get_buffer: unknown buffer");
    }
    if (!p) return c.fail(RTRT_ERR_STATE, "rtrt_get_buffer: buffer not allocated yet (no mesh set / no frame rendered?)");
    *dptr = p; *bytes = b;
    return RTRT_OK;
}

int rtrt_read_buffer(rtrt_ctx* ctx, int which, void* dst, size_t dst_bytes) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    void* p; size_t b;
    int rc = rtrt_get_buffer(ctx, which, &p, &b);
    if (rc) return rc;
    if (!dst || dst_bytes < b) return c.fail(RTRT_ERR_ARG, "rtrt_read_buffer: destination too small");
    RTRT_CUDA(c, cudaSetDevice(c.device));
    RTRT_CUDA(c, cudaMemcpyAsync(dst, p, b, cudaMemcpyDeviceToHost, c.stream));
    RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    return RTRT_OK;
}

// Frames in flight: D2D snapshot on the main stream (ordered after the kernels that produce the buffer), D2H of the snapshot
// on the copy stream, so the next frame's kernels overlap the transfer.
int rtrt_read_buffer_async(rtrt_ctx* ctx, int which, void* dst, size_t dst_bytes) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    void* p; size_t b;
    int rc = rtrt_get_buffer(ctx, which, &p, &b);
    if (rc) return rc;
    if (!dst || dst_bytes < b) return c.fail(RTRT_ERR_ARG, "rtrt_read_buffer_async: destination too small");
    RTRT_CUDA(c, cudaSetDevice(c.device));
    if (c.stageBytes < b) {
        RTRT_CUDA(c, cudaStreamSynchronize(c.copyStream));
        dfree(c.dStage); c.stageBytes = 0;
        RTRT_CUDA(c, cudaMalloc(&c.dStage, b));
        c.stageBytes = b;
    }
    if (c.readPending) RTRT_CUDA(c, cudaStreamWaitEvent(c.stream, c.evReadDone, 0));      // the previous read-back still owns the snapshot
    RTRT_CUDA(c, cudaMemcpyAsync(c.dStage, p, b, cudaMemcpyDeviceToDevice, c.stream));
    RTRT_CUDA(c, cudaEventRecord(c.evStaged, c.stream));
    RTRT_CUDA(c, cudaStreamWaitEvent(c.copyStream, c.evStaged, 0));
    RTRT_CUDA(c, cudaMemcpyAsync(dst, c.dStage, b, cudaMemcpyDeviceToHost, c.copyStream));
    RTRT_CUDA(c, cudaEventRecord(c.evReadDone, c.copyStream));
    c.readPending = true;
    return RTRT_OK;
}

// Main stream waits for everything queued on the copy stream (so an event recorded on rtrt_get_stream afterwards covers it).
int rtrt_join(rtrt_ctx* ctx) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    RTRT_CUDA(c, cudaSetDevice(c.device));
    if (c.readPending) RTRT_CUDA(c, cudaStreamWaitEvent(c.stream, c.evReadDone, 0));
    if (c.geomPending) RTRT_CUDA(c, cudaStreamWaitEvent(c.stream, c.evGeomReady, 0));
    return RTRT_OK;
}

int rtrt_get_counts(rtrt_ctx* ctx, uint32_t* nt, uint32_t* nb, uint32_t* nbp) {
    CTX_OR_FAIL(ctx);
    if (nt) *nt = ctx->c.numTris;
    if (nb) *nb = ctx->c.numBatches;
    if (nbp) *nbp = ctx->c.numBatchesPadded;
    return RTRT_OK;
}

int rtrt_sync(rtrt_ctx* ctx) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    RTRT_CUDA(c, cudaSetDevice(c.device));
    RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    RTRT_CUDA(c, cudaStreamSynchronize(c.copyStream));
    RTRT_CUDA(c, cudaStreamSynchronize(c.uploadStream));
    c.readPending = false;
    { int rc = check_wide_error(c); if (rc) return rc; }
    return check_build_error(c);
}

int rtrt_get_timings(rtrt_ctx* ctx, rtrt_timings* out) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (!out) return c.fail(RTRT_ERR_ARG, "rtrt_get_timings: null out");
    RTRT_CUDA(c, cudaSetDevice(c.device));
    RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    { int rc = check_build_error(c); if (rc) return rc; }
    { int rc = check_wide_error(c); if (rc) return rc; }
    float* slots[4] = {&c.timings.build_bvh_ms, &c.timings.traverse_ms, &c.timings.pathtrace_ms, &c.timings.denoise_ms};
    for (int s = 0; s < 4; ++s)
        if (c.evPending[s]) {
            float ms = 0.0f;
            *slots[s] = (cudaEventElapsedTime(&ms, c.ev[s][0], c.ev[s][1]) == cudaSuccess) ? ms : 0.0f;
            if (s == STAGE_BUILD) {
                c.timings.wide_build_ms = 0.0f;
                if (c.evWidePending) {                                    // split the stage: LBVH build | collapse to the wide layout
                    float lb = 0.0f, wb = 0.0f;
                    if (cudaEventElapsedTime(&lb, c.ev[s][0], c.evWideBegin) == cudaSuccess && cudaEventElapsedTime(&wb, c.evWideBegin, c.ev[s][1]) == cudaSuccess) {
                        c.timings.build_bvh_ms = lb; c.timings.wide_build_ms = wb;
                    }
                    c.evWidePending = false;
                }
            }
            c.evPending[s] = false;
        }
    if (c.raysOnDevice && c.dRayCounter) {
        unsigned long long r[5] = {0, 0, 0, 0, 0};
        RTRT_CUDA(c, cudaMemcpy(r, c.dRayCounter, sizeof(r), cudaMemcpyDeviceToHost));
        c.timings.rays_traced = r[0]; c.timings.node_visits = r[1]; c.timings.leaf_tests = r[2]; c.raysOnDevice = false;
        c.timings.rays_traced_total = r[4];
    }
    if (c.kevPending) {
        for (int k = 0; k < 6; ++k) {
            float ms = 0.0f;
            c.timings.pt_kernel_ms[k] = (cudaEventElapsedTime(&ms, c.kev[k], c.kev[k + 1]) == cudaSuccess) ? ms : 0.0f;
        }
        c.kevPending = false;
    }
    if (c.countVisits && c.dWf) {                        // per-round diagnostics of the wavefront counters (pathtrace_wf.cu)
        uint32_t w[64];
        RTRT_CUDA(c, cudaMemcpy(w, c.dWf, sizeof(w), cudaMemcpyDeviceToHost));
        for (int r = 0; r < 4; ++r) { c.timings.round_rays[r] = w[r]; c.timings.round_node_visits[r] = w[56 + r]; c.timings.round_leaf_tests[r] = w[60 + r]; }
    }
    c.timings.kernel_launches = c.launches;
    *out = c.timings;
    return RTRT_OK;
}

void* rtrt_get_stream(rtrt_ctx* ctx) { return ctx ? (void*)ctx->c.stream : nullptr; }

// ---------------------------------------------------------------------------------- frame inputs
int rtrt_set_camera(rtrt_ctx* ctx, const void* blob, size_t bytes, int reset_history) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (!blob || bytes != sizeof(Camera)) return c.fail(RTRT_ERR_ARG, "rtrt_set_camera: blob must be sizeof(Camera) = 176 bytes");
    std::memcpy(&c.camera, blob, sizeof(Camera));
    if (reset_history || !c.cameraSet) c.historyCamera = history_setup(c.camera);      // src/kernel.cu:131-135
    c.cameraSet = true;
    return RTRT_OK;
}

int rtrt_load_camera_file(rtrt_ctx* ctx, const char* path) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    Camera cam;
    FILE* f = path ? fopen(path, "rb") : nullptr;
    if (!f) return c.fail(RTRT_ERR_IO, "rtrt_load_camera_file: cannot open file");
    const size_t n = fread(&cam, 1, sizeof(cam), f);
    fclose(f);
    if (n != sizeof(cam)) return c.fail(RTRT_ERR_IO, "rtrt_load_camera_file: short file");
    return rtrt_set_camera(ctx, &cam, sizeof(cam), 1);
}

int rtrt_save_camera_file(rtrt_ctx* ctx, const char* path) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (!c.cameraSet) return c.fail(RTRT_ERR_STATE, "rtrt_save_camera_file: no camera set");
    FILE* f = path ? fopen(path, "wb") : nullptr;
    if (!f) return c.fail(RTRT_ERR_IO, "rtrt_save_camera_file: cannot open file");
    const size_t n = fwrite(&c.camera, 1, sizeof(Camera), f);
    fclose(f);
    return n == sizeof(Camera) ? RTRT_OK : c.fail(RTRT_ERR_IO, "rtrt_save_camera_file: short write");
}

int rtrt_set_render_size(rtrt_ctx* ctx, int32_t w, int32_t h) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (w < 8 || h < 8 || w > 16384 || h > 16384) return c.fail(RTRT_ERR_ARG, "rtrt_set_render_size: size out of range");
    c.width = w; c.height = h; c.rowBegin = 0; c.rowEnd = h;
    return RTRT_OK;
}

// src/kernel.cu:78-101 (int *= float truncates; clampi; height follows the width at 16:9)
int rtrt_dynamic_resolution_step(const rtrt_config* cfg, int32_t curW, int32_t curH, float deltaTime, int32_t* outW, int32_t* outH) {
    if (!cfg || !outW || !outH || curW <= 0 || !(deltaTime > 0.0f) || !(cfg->targetFps > 2.0f)) return RTRT_ERR_ARG;
    if (!cfg->useDynamicResolution) { *outW = curW; *outH = curH; return RTRT_OK; }
    int renderWidth = curW;
    const float targetFrameTimeHigh = 1000.0f / (cfg->targetFps - 2), targetFrameTimeLow = 1000.0f / (cfg->targetFps + 2);
    if (targetFrameTimeHigh < deltaTime || targetFrameTimeLow > deltaTime) {
        float ratio = (1000.0f / cfg->targetFps) / deltaTime;
        ratio = sqrtf(ratio);
        renderWidth = (int)((float)renderWidth * ratio);
    }
    renderWidth = renderWidth + ((renderWidth % 16 < 8) ? (-renderWidth % 16) : (16 - renderWidth % 16));
    renderWidth = renderWidth < cfg->minWidth ? cfg->minWidth : (renderWidth > cfg->maxWidth ? cfg->maxWidth : renderWidth);
    *outW = renderWidth; *outH = (renderWidth / 16) * 9;
    return RTRT_OK;
}

int rtrt_set_frame_number(rtrt_ctx* ctx, int32_t n) { CTX_OR_FAIL(ctx); ctx->c.frameNum = n; return RTRT_OK; }

int rtrt_set_sky_params(rtrt_ctx* ctx, const rtrt_sky_params* p) {
    CTX_OR_FAIL(ctx);
    if (p) ctx->c.skyParams = *p; else rtrt_set_sky_params_defaults(&ctx->c.skyParams);
    ctx->c.skyDirty = true;
    return RTRT_OK;
}

int rtrt_get_sun(rtrt_ctx* ctx, float* d3, float* cosMax) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (c.skyDirty || !c.dSky) { int rc = ensure_frame_resources(c); if (rc) return rc; }
    if (d3) { d3[0] = c.sunDir.x; d3[1] = c.sunDir.y; d3[2] = c.sunDir.z; }
    if (cosMax) *cosMax = c.sunAngleCosThetaMax;
    return RTRT_OK;
}

int rtrt_set_denoise_params(rtrt_ctx* ctx, const rtrt_denoise_params* p, const rtrt_pass_settings* s) {
    CTX_OR_FAIL(ctx);
    rtrt_denoise_params d; rtrt_pass_settings ps;
    rtrt_set_denoise_defaults(&d, &ps);
    ctx->c.denoiseParams = p ? *p : d;
    ctx->c.passSettings = s ? *s : ps;
    return RTRT_OK;
}

int rtrt_set_materials(rtrt_ctx* ctx, const rtrt_material* m, int32_t nm, const int32_t* idx, uint32_t nidx) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (m && nm <= 0) return c.fail(RTRT_ERR_ARG, "rtrt_set_materials: bad count");
    static_assert(sizeof(rtrt_material) == sizeof(SurfaceMaterial), "material layout");
    c.materialsUser = (m != nullptr) || (idx != nullptr);
    return set_materials(c, m, nm, idx, (int)nidx);
}

int rtrt_set_texture(rtrt_ctx* ctx, int32_t slot, int32_t w, int32_t h, const uint16_t* rgba) {
    CTX_OR_FAIL(ctx);
    return set_texture(ctx->c, slot, w, h, rgba);
}

int rtrt_set_row_range(rtrt_ctx* ctx, int32_t b, int32_t e) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (b < 0 || e <= b || e > c.height) return c.fail(RTRT_ERR_ARG, "rtrt_set_row_range: bad range");
    c.rowBegin = b; c.rowEnd = e;
    return RTRT_OK;
}

int rtrt_set_row_strips(rtrt_ctx* ctx, int32_t strip_height, int32_t num_shards, int32_t shard) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (num_shards <= 1) { c.stripH = 0; c.stripShards = 1; c.stripShard = 0; return RTRT_OK; }
    if (strip_height < 1 || shard < 0 || shard >= num_shards) return c.fail(RTRT_ERR_ARG, "rtrt_set_row_strips: bad arguments");
    c.stripH = strip_height; c.stripShards = num_shards; c.stripShard = shard;
    return RTRT_OK;
}

int rtrt_set_data_dir(rtrt_ctx* ctx, const char* dir) {
    CTX_OR_FAIL(ctx);
    if (!dir) return ctx->c.fail(RTRT_ERR_ARG, "rtrt_set_data_dir: null");
    ctx->c.dataDir = dir;
    return RTRT_OK;
}

// ---------------------------------------------------------------------------------- stages 2 and 3
int rtrt_pathtrace(rtrt_ctx* ctx) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    RTRT_CUDA(c, cudaSetDevice(c.device));
    RTRT_CUDA(c, cudaEventRecord(c.ev[STAGE_PATHTRACE][0], c.stream));
    int rc = launch_pathtrace(c);
    if (rc) return rc;
    RTRT_CUDA(c, cudaEventRecord(c.ev[STAGE_PATHTRACE][1], c.stream));
    c.evPending[STAGE_PATHTRACE] = true; c.raysOnDevice = true;
    c.historyCamera = history_setup(c.camera);                         // src/kernel.cu:357
    return RTRT_OK;
}

int rtrt_denoise(rtrt_ctx* ctx) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    RTRT_CUDA(c, cudaSetDevice(c.device));
    RTRT_CUDA(c, cudaEventRecord(c.ev[STAGE_DENOISE][0], c.stream));
    int rc = launch_denoise(c);
    if (rc) return rc;
    RTRT_CUDA(c, cudaEventRecord(c.ev[STAGE_DENOISE][1], c.stream));
    c.evPending[STAGE_DENOISE] = true;
    return RTRT_OK;
}

int rtrt_set_postprocess_params(rtrt_ctx* ctx, const rtrt_postprocess_params* p) {
    CTX_OR_FAIL(ctx);
    if (p) ctx->c.postParams = *p; else rtrt_postprocess_defaults(&ctx->c.postParams);
    if (!(ctx->c.postParams.gamma > 0.0f)) { rtrt_postprocess_defaults(&ctx->c.postParams); return ctx->c.fail(RTRT_ERR_ARG, "rtrt_set_postprocess_params: gamma must be > 0"); }
    return RTRT_OK;
}

int rtrt_set_output_size(rtrt_ctx* ctx, int32_t w, int32_t h) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (w < 8 || h < 8 || w > 16384 || h > 16384) return c.fail(RTRT_ERR_ARG, "rtrt_set_output_size: size out of range");
    c.outW = w; c.outH = h;
    return RTRT_OK;
}

int rtrt_postprocess(rtrt_ctx* ctx, float delta_time_ms) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    RTRT_CUDA(c, cudaSetDevice(c.device));
    return launch_postprocess(c, delta_time_ms);
}

int rtrt_write_ppm(rtrt_ctx* ctx, const char* path) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (!path) return c.fail(RTRT_ERR_ARG, "rtrt_write_ppm: null path");
    if (!c.dOutRgba) return c.fail(RTRT_ERR_STATE, "rtrt_write_ppm: no post-processed frame (call rtrt_postprocess)");
    const size_t n = (size_t)c.ppOutW * c.ppOutH;
    std::vector<unsigned char> rgba(n * 4), rgb(n * 3);
    RTRT_CUDA(c, cudaMemcpyAsync(rgba.data(), c.dOutRgba, n * 4, cudaMemcpyDeviceToHost, c.stream));
    RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    for (size_t i = 0; i < n; ++i) { rgb[3 * i] = rgba[4 * i]; rgb[3 * i + 1] = rgba[4 * i + 1]; rgb[3 * i + 2] = rgba[4 * i + 2]; }
    FILE* f = fopen(path, "wb");
    if (!f) return c.fail(RTRT_ERR_IO, "rtrt_write_ppm: cannot open file");
    fprintf(f, "P6\n%d %d\n255\n", c.ppOutW, c.ppOutH);
    const bool ok = fwrite(rgb.data(), 1, rgb.size(), f) == rgb.size();
    fclose(f);
    return ok ? RTRT_OK : c.fail(RTRT_ERR_IO, "rtrt_write_ppm: short write");
}

int rtrt_accumulate(rtrt_ctx* ctx, int reset) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    RTRT_CUDA(c, cudaSetDevice(c.device));
    return launch_accumulate(c, reset);
}

int rtrt_render_frame(rtrt_ctx* ctx) {
    CTX_OR_FAIL(ctx);
    int rc = rtrt_build_bvh(ctx);            // src/kernel.cu:330-331
    if (rc) return rc;
    rc = rtrt_pathtrace(ctx);                // :336-357
    if (rc) return rc;
    rc = rtrt_denoise(ctx);                  // :360
    if (rc) return rc;
    ctx->c.frameNum++;                       // UpdateFrame of the next draw (src/kernel.cu:61-63)
    return RTRT_OK;
}

}  // extern "C"
