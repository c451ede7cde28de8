// Internal context behind the C ABI (include/rtrt.h).  Plays the role of the reference's RayTracer
// object (src/kernel.cuh:431-621): owns every device allocation of the hot path, one CUDA stream,
// and the per-stage CUDA-event timers.  One context per GPU; not thread-safe per context.
#pragma once
#include "../../include/rtrt.h"
#include "rtrt_types.h"
#include "shading.cuh"
#include <cuda_runtime.h>
#include <string>
#include <cstdio>

namespace rtrt {

struct Context {
    int device = 0;
    int smCount = 148;                       // SMs of `device` (rtrt_create)
    bool buildAttrSet = false;               // k_build_blas shared-memory opt-in done on this device
    int traceResident = 0, traceResidentWide = 0;   // resident k_pt_trace CTAs on this device (reference layout / wide layout)
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
    int mortonBits = 60;                                  // option "morton_bits": 60 = the reference tool's code (20 bits / axis), 63 = 21 bits / axis
    bool globalSort = false;                              // option "global_sort": device Morton pre-sort of the triangles before batching
    uint32_t* dIdxSorted = nullptr; uint32_t* dTriPerm = nullptr; void* dPresort = nullptr; size_t presortBytes = 0;
    bool bvhValid = false;
    // ---- wide traversal layout (option wide_bvh; wide_build.cu) ----
    bool wideBvh = false, wideValid = false, wideStatPending = false;
    // wide_bvh = 2: the collapse to the wide layout runs on a side stream, concurrently with the primary-ray kernels (which then do not need it)
    cudaStream_t wideStream = nullptr; cudaEvent_t evWideFork = nullptr, evWideDone = nullptr; bool widePending = false;
    cudaStream_t wideTlasStream = nullptr; cudaEvent_t evTlasFork = nullptr, evTlasDone = nullptr;   // the top-level collapse runs beside the batches' collapse
    bool widePrimary = true;                              // option wide_bvh = 2: primary rays keep the reference layout (coherent rays: its two-child nodes hit L1), bounce rays the wide one
    void* dWideNodes = nullptr; void* dWideTris = nullptr; void* dWideQueue = nullptr; size_t wideNodeCap = 0;
    uint32_t wideRoot = 0, wideNodes = 0;
    uint32_t tlasSpinLimit = 1u << 22;                    // option "tlas_spin_limit" (test knob: 0 forces the time-out path)
    bool keepAllTris = false;                             // option "keep_all_triangles"
    bool buildErrPending = false;                         // a fused build was launched since the error word was last read

    // ---- traversal scratch for the host-buffer entry point ----
    Ray* dRayStage = nullptr; Hit* dHitStage = nullptr; uint32_t rayStageCap = 0;
    int stackSize = 16;      // src/traverse.h:28
    int maxLoop = 1024;      // src/traverse.h:132
    bool countVisits = false;
    bool wavefront = true;                 // path-trace driver: wavefront queues (default) or the single megakernel
    void* dWf = nullptr; size_t wfBytes = 0;   // wavefront path state + queues (pathtrace_wf.cu)
    bool anyHitRays = true;                    // option any_hit_rays: shadow / last-segment rays stop at the first accepted triangle
    int ptChunks = 1;                          // option pt_chunks: independent pixel chunks of one frame, pipelined on side streams
    static constexpr int kMaxChunks = 8;
    cudaStream_t chunkStream[kMaxChunks] = {}; cudaEvent_t chunkFork = nullptr; cudaEvent_t chunkJoin[kMaxChunks] = {};

    // ---- frame state: ConstBuffer (src/kernel.cuh:211-231) + Buffer2DManager (src/init.cu:441-584) ----
    int width = 0, height = 0;               // render size
    int rowBegin = 0, rowEnd = 0;            // rows this GPU traces / denoises (multi-GPU row bands); default = whole frame
    int stripH = 0, stripShards = 1, stripShard = 0;   // rtrt_set_row_strips: path trace / accumulate rows with (row / stripH) % stripShards == stripShard
    int frameNum = 1;                        // src/kernel.cu:61-63 (first frame is 1)
    bool cameraSet = false;
    Camera camera{}; Camera prevCamera{}; HistoryCamera historyCamera{};
    rtrt_sky_params skyParams{}; bool skyDirty = true;
    Float3 sunDir{0, 1, 0}; float sunAngleCosThetaMax = 1.0f;
    std::string dataDir;
    float* dHosek = nullptr;                 // Hosek-Wilkie coefficient sets (2460 floats)
    uint8_t* dBlueNoise = nullptr;           // sobol | scramble | rank
    float* dSky = nullptr; float* dSkyPdf = nullptr; float* dSkyCdf = nullptr;
    float* dSun = nullptr; float* dSunPdf = nullptr; float* dSunCdf = nullptr;
    SurfaceMaterial* dMaterials = nullptr; int* dMaterialsIdx = nullptr; int numMaterials = 0, numMaterialsIdx = 0;
    bool materialsUser = false;
    MipTex tex[2]{}; void* dTexMem[2] = {nullptr, nullptr};
    // G-buffers, linear, reference formats (src/init.cu:473-500)
    void* dColor = nullptr; void* dNormal = nullptr; void* dAlbedo = nullptr; void* dDepth = nullptr; void* dMotion = nullptr;
    void* dAccum = nullptr; void* dHistoryColor = nullptr; void* dHistoryDepth = nullptr; void* dHistoryAlbedo = nullptr;
    void* dNoise = nullptr; void* dNoise16 = nullptr; void* dColorTmp = nullptr; void* dGuide = nullptr;
    int gbufW = 0, gbufH = 0;                 // layout (row stride) of the current frame's buffers
    int gbufCapW = 0, gbufCapH = 0;           // allocated capacity; a smaller render size re-uses the storage and keeps the history
    int accumW = 0, accumH = 0, histW = 0, histH = 0;   // sizes the two history images were written at (historyDim, src/kernel.cu:83-84,266)
    void* dAccumRadiance = nullptr; size_t accumRadiancePx = 0;      // progressive fp32 accumulation (rtrt_accumulate)
    // post-processing (postprocess.cu): output (screen) size, RGBA8 frame, 1/64 colour, exposure[4] + histogram[64]
    int outW = 0, outH = 0, ppOutW = 0, ppOutH = 0, ppW64 = 0, ppH64 = 0;
    void* dOutRgba = nullptr; void* dColor64 = nullptr; void* dExposure = nullptr; size_t ppCap64 = 0;
    rtrt_postprocess_params postParams{};
    unsigned long long* dRayCounter = nullptr;
    rtrt_denoise_params denoiseParams{}; rtrt_pass_settings passSettings{};

    // ---- copy engine (frames in flight): a second stream carries the host<->device copies of rtrt_update_vertices and
    // rtrt_read_buffer_async so that they overlap the kernels of the neighbouring frames; ordering is by events only.
    cudaStream_t copyStream = nullptr, uploadStream = nullptr;      // read-backs / uploads (PCIe is full duplex: one in-order stream each)
    cudaEvent_t evGeomReady = nullptr, evGeomFree = nullptr, evStaged = nullptr, evReadDone = nullptr;
    bool geomPending = false, geomFreeValid = false, readPending = false;
    void* dStage = nullptr; size_t stageBytes = 0;       // device staging copy of the buffer being read back

    // ---- multi-GPU (multi.cu): the group this context is a rank of, and the denoiser's halo hook ----
    void* multi = nullptr;                               // rtrt::Multi
    int bandHaloRows = -1;                               // rows around the band that rtrt_denoise recomputes: -1 = the full dependency halo (80 / 112 rows,
                                                         // nothing is exchanged between passes), 0 = none (the halo rows are exchanged: haloFn)
    int (*haloFn)(Context&, void* colour, int rows) = nullptr;   // called before every pass that reads `rows` neighbouring rows of `colour`

    // ---- bookkeeping ----
    unsigned long long launches = 0;                     // kernels launched by this library
    cudaEvent_t ev[4][2] = {};                           // [build, traverse, pathtrace, denoise][begin, end]
    cudaEvent_t evWideBegin = nullptr; bool evWidePending = false;   // between the LBVH build and the wide collapse of rtrt_build_bvh
    bool evPending[4] = {false, false, false, false};
    bool kernelTiming = false; cudaEvent_t kev[7] = {}; bool kevPending = false;   // option kernel_timing: boundaries of the path-trace kernels
    bool raysOnDevice = false;                           // rays_traced of the last pathtrace still in dRayCounter
    rtrt_timings timings{};

    int fail(int code, const char* msg) { lastError = msg; return code; }
    int failCuda(cudaError_t e, const char* what, const char* file, int line) {
        char buf[512];
        snprintf(buf, sizeof(buf), "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
        lastError = buf;
        return RTRT_ERR_CUDA;
    }
};

enum { STAGE_BUILD = 0, STAGE_TRAVERSE = 1, STAGE_PATHTRACE = 2, STAGE_DENOISE = 3 };

#define RTRT_CUDA(ctx, call)                                                         \
    do {                                                                             \
        cudaError_t _e = (call);                                                     \
        if (_e != cudaSuccess) return (ctx).failCuda(_e, #call, __FILE__, __LINE__); \
    } while (0)

// stage launchers (each enqueues on ctx.stream, no host sync)
int launch_build_bvh(Context& c);
int launch_build_tlas_global(Context& c);
int launch_presort(Context& c);             // bvh_presort.cu
int ensure_wide(Context& c);                // traverse.cu: builds the wide layout if option wide_bvh is on and it is stale
int launch_wide_build(Context& c);          // wide_build.cu: 8-wide compressed nodes + 48-byte triangles from the arrays of the build
int check_wide_error(Context& c);           // wide_build.cu: after a stream sync
int wide_join(Context& c);                  // wide_build.cu: main stream waits for a wide build running on the side stream
int wait_geometry(Context& c);              // capi.cu: main stream waits for a pending rtrt_update_vertices upload
int check_build_error(Context& c);           // capi.cu: after a stream sync, turns the fused build's time-out word into RTRT_ERR_CUDA
int geometry_consumed(Context& c);          // capi.cu: records that the main stream has read the vertex buffers
int launch_traverse(Context& c, const Ray* dRays, uint32_t n, Hit* dHits, int anyHit, unsigned long long* dCounters);
int launch_pathtrace(Context& c);
int launch_denoise(Context& c);
int launch_accumulate(Context& c, int reset);
int launch_postprocess(Context& c, float deltaTimeMs);   // postprocess.cu
void multi_release(Context& c);             // multi.cu: leaves the multi-GPU group (rtrt_destroy)
int local_rows(const Context& c);            // rows of the frame this context path-traces (band or strips)
int ensure_frame_resources(Context& c);      // frame.cu: G-buffers, tables, default materials/textures, sky regeneration
void free_frame_resources(Context& c);
int set_texture(Context& c, int slot, int w, int h, const uint16_t* rgba);
int set_materials(Context& c, const void* materials, int numMaterials, const int* materialsIdx, int numIdx);

}  // namespace rtrt

// the opaque handle of include/rtrt.h
struct rtrt_ctx { rtrt::Context c; };
