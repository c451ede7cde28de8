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

struct rtrt_ctx { Context c; };

namespace rtrt {

template <typename T>
static cudaError_t dalloc(T*& p, size_t count) {
    p = nullptr;
    return cudaMalloc((void**)&p, count ? count * sizeof(T) : sizeof(T));
}
template <typename T>
static void dfree(T*& p) { if (p) cudaFree(p); p = nullptr; }

static void free_mesh(Context& c) {
    if (c.ownsMesh) { dfree(c.dPos); dfree(c.dNrm); dfree(c.dIdx); }
    c.dPos = nullptr; c.dNrm = nullptr; c.dIdx = nullptr; c.ownsMesh = false;
    dfree(c.dTriangles); dfree(c.dAabbs); dfree(c.dMorton); dfree(c.dReorderIdx); dfree(c.dNodes); dfree(c.dMortonPre);
    dfree(c.dTlasAabbs); dfree(c.dTlasMorton); dfree(c.dTlasReorderIdx); dfree(c.dTlasNodes); dfree(c.dTlasMortonPre);
    if (c.dSortTmp) { cudaFree(c.dSortTmp); c.dSortTmp = nullptr; c.sortTmpBytes = 0; }
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
    if (cudaSetDevice(device) != cudaSuccess) { delete h; return RTRT_ERR_CUDA; }
    if (cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking) != cudaSuccess) { delete h; return RTRT_ERR_CUDA; }
    cudaEventCreate(&c.evA); cudaEventCreate(&c.evB);
    if (cudaMalloc((void**)&c.dTicket, 64) != cudaSuccess) { delete h; return RTRT_ERR_CUDA; }
    *out = h;
    return RTRT_OK;
}

int rtrt_destroy(rtrt_ctx* ctx) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    cudaSetDevice(c.device);
    cudaStreamSynchronize(c.stream);
    free_mesh(c);
    dfree(c.dRayStage); dfree(c.dHitStage); dfree(c.dTicket);
    if (c.evA) cudaEventDestroy(c.evA);
    if (c.evB) cudaEventDestroy(c.evB);
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

int rtrt_set_mesh_bin(rtrt_ctx* ctx, const char* path) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (!path) return c.fail(RTRT_ERR_ARG, "rtrt_set_mesh_bin: null path");
    FILE* f = fopen(path, "rb");
    if (!f) return c.fail(RTRT_ERR_IO, "rtrt_set_mesh_bin: cannot open file");
    uint32_t n = 0;
    if (fread(&n, sizeof(n), 1, f) != 1 || n == 0) { fclose(f); return c.fail(RTRT_ERR_IO, "rtrt_set_mesh_bin: bad header"); }
    std::vector<Triangle> tris(n);
    if (fread(tris.data(), sizeof(Triangle), n, f) != n) { fclose(f); return c.fail(RTRT_ERR_IO, "rtrt_set_mesh_bin: truncated file (legacy 64-byte layout?)"); }
    fclose(f);
    // un-indexed triangle soup -> indexed buffers (3 vertices per triangle)
    std::vector<float> pos((size_t)9 * n), nrm((size_t)9 * n);
    std::vector<uint32_t> idx((size_t)3 * n);
    for (uint32_t i = 0; i < n; ++i) {
        const Float3* v[3] = {&tris[i].v1, &tris[i].v2, &tris[i].v3};
        const Float3* m[3] = {&tris[i].n1, &tris[i].n2, &tris[i].n3};
        for (int k = 0; k < 3; ++k) {
            std::memcpy(&pos[(size_t)9 * i + 3 * k], v[k], 12);
            std::memcpy(&nrm[(size_t)9 * i + 3 * k], m[k], 12);
            idx[(size_t)3 * i + k] = 3 * i + k;
        }
    }
    return rtrt_set_mesh(ctx, pos.data(), nrm.data(), idx.data(), 3 * n, n);
}

int rtrt_set_option_i(rtrt_ctx* ctx, const char* name, int64_t value) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (!name) return c.fail(RTRT_ERR_ARG, "rtrt_set_option_i: null name");
    if (!strcmp(name, "keep_presort_morton")) { c.keepPresortMorton = value != 0; return RTRT_OK; }
    if (!strcmp(name, "stack_size")) { if (value < 1 || value > 32) return c.fail(RTRT_ERR_ARG, "stack_size must be 1..32"); c.stackSize = (int)value; return RTRT_OK; }
    if (!strcmp(name, "max_loop")) { if (value < 1) return c.fail(RTRT_ERR_ARG, "max_loop must be >= 1"); c.maxLoop = (int)value; return RTRT_OK; }
    return c.fail(RTRT_ERR_ARG, "rtrt_set_option_i: unknown option");
}

int rtrt_build_bvh(rtrt_ctx* ctx) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    RTRT_CUDA(c, cudaSetDevice(c.device));
    RTRT_CUDA(c, cudaEventRecord(c.evA, c.stream));
    int rc = launch_build_bvh(c);
    if (rc) return rc;
    RTRT_CUDA(c, cudaEventRecord(c.evB, c.stream));
    c.bvhValid = true;
    c.timings.build_bvh_ms = -1.0f;     // resolved lazily in rtrt_get_timings
    return RTRT_OK;
}

int rtrt_traverse(rtrt_ctx* ctx, const rtrt_ray* d_rays, uint32_t n, rtrt_hit* d_hits, int any_hit) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if ((!d_rays || !d_hits) && n) return c.fail(RTRT_ERR_ARG, "rtrt_traverse: null buffer");
    RTRT_CUDA(c, cudaSetDevice(c.device));
    c.timings.rays_traced = n;
    return launch_traverse(c, reinterpret_cast<const Ray*>(d_rays), n, reinterpret_cast<Hit*>(d_hits), any_hit, nullptr);
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
    c.timings.rays_traced = n;
    return RTRT_OK;
}

int rtrt_get_buffer(rtrt_ctx* ctx, int which, void** dptr, size_t* bytes) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (!dptr || !bytes) return c.fail(RTRT_ERR_ARG, "rtrt_get_buffer: null out pointer");
    const size_t n = (size_t)c.numBatches * kBatchSize, bp = c.numBatchesPadded;
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
        default: return c.fail(RTRT_ERR_ARG, "rtrt_get_buffer: unknown buffer");
    }
    if (!p) return c.fail(RTRT_ERR_STATE, "rtrt_get_buffer: buffer not allocated (no mesh set?)");
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
    return RTRT_OK;
}

int rtrt_get_timings(rtrt_ctx* ctx, rtrt_timings* out) {
    CTX_OR_FAIL(ctx);
    Context& c = ctx->c;
    if (!out) return c.fail(RTRT_ERR_ARG, "rtrt_get_timings: null out");
    RTRT_CUDA(c, cudaSetDevice(c.device));
    RTRT_CUDA(c, cudaStreamSynchronize(c.stream));
    if (c.timings.build_bvh_ms < 0.0f) {
        float ms = 0.0f;
        if (cudaEventElapsedTime(&ms, c.evA, c.evB) == cudaSuccess) c.timings.build_bvh_ms = ms; else c.timings.build_bvh_ms = 0.0f;
    }
    c.timings.kernel_launches = c.launches;
    *out = c.timings;
    return RTRT_OK;
}

void* rtrt_get_stream(rtrt_ctx* ctx) { return ctx ? (void*)ctx->c.stream : nullptr; }

}  // extern "C"
