// Internal context behind the C ABI (include/rtrt.h).  Plays the role of the reference's RayTracer
// object (src/kernel.cuh:431-621): owns every device allocation of the hot path, one CUDA stream,
// and the per-stage CUDA-event timers.  One context per GPU; not thread-safe per context.
#pragma once
#include "../../include/rtrt.h"
#include "rtrt_types.h"
#include <cuda_runtime.h>
#include <string>
#include <cstdio>

namespace rtrt {

struct Context {
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string lastError;
    rtrt_config cfg{};

    // ---- mesh inputs (src/init.cu:139-150: indexBuffer, vertexBuffer, normalBuffer) ----
    float* dPos = nullptr; float* dNrm = nullptr; uint32_t* dIdx = nullptr;
    bool ownsMesh = false;
    uint32_t numVerts = 0, numTris = 0, numBatches = 0, numBatchesPadded = 0;

    // ---- stage 1 outputs (src/init.cu:159-187) ----
    Triangle* dTriangles = nullptr; AABB* dAabbs = nullptr; uint32_t* dMorton = nullptr; uint32_t* dReorderIdx = nullptr;
    BVHNode* dNodes = nullptr; uint32_t* dMortonPre = nullptr;
    AABB* dTlasAabbs = nullptr; uint32_t* dTlasMorton = nullptr; uint32_t* dTlasReorderIdx = nullptr;
    BVHNode* dTlasNodes = nullptr; uint32_t* dTlasMortonPre = nullptr;
    uint32_t* dTicket = nullptr;
    void* dSortTmp = nullptr; size_t sortTmpBytes = 0;     // scratch of the global (multi-CTA) paths
    bool keepPresortMorton = false;
    bool bvhValid = false;

    // ---- traversal scratch for the host-buffer entry point ----
    Ray* dRayStage = nullptr; Hit* dHitStage = nullptr; uint32_t rayStageCap = 0;
    int stackSize = 16;      // src/traverse.h:28
    int maxLoop = 1024;      // src/traverse.h:132

    // ---- bookkeeping ----
    unsigned long long launches = 0;                     // kernels launched by this library
    cudaEvent_t evA = nullptr, evB = nullptr;
    rtrt_timings timings{};

    int fail(int code, const char* msg) { lastError = msg; return code; }
    int failCuda(cudaError_t e, const char* what, const char* file, int line) {
        char buf[512];
        snprintf(buf, sizeof(buf), "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
        lastError = buf;
        return RTRT_ERR_CUDA;
    }
};

#define RTRT_CUDA(ctx, call)                                                         \
    do {                                                                             \
        cudaError_t _e = (call);                                                     \
        if (_e != cudaSuccess) return (ctx).failCuda(_e, #call, __FILE__, __LINE__); \
    } while (0)

// stage launchers (each enqueues on ctx.stream, no host sync)
int launch_build_bvh(Context& c);
int launch_build_tlas_global(Context& c);
int launch_traverse(Context& c, const Ray* dRays, uint32_t n, Hit* dHits, int anyHit, unsigned long long* dCounters);

}  // namespace rtrt
