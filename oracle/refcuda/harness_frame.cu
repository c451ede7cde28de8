// oracle/refcuda/harness_frame.cu — extern "C" launchers around the REFERENCE'S OWN stage-2 / stage-3 / sky kernels
// (test / baseline infrastructure).  The kernel bodies come from /root/reference/src/{pathtrace,temporalDenoising,sky,scan}.cuh
// and everything they include (never copied into this repo); this file only allocates the surfaces the way
// Buffer2DManager::init does (src/init.cu:441-584, src/kernel.cuh:344-410) and reproduces the launch sequences of
// RayTracer::draw (src/kernel.cu:283-354) and RayTracer::TemporalSpatialDenoising (src/denoising.cu:48-186).
// Built by oracle/refcuda/build_ref.py into oracle/_ref/ together with harness.cu.
#include "kernel_cu.inc"          /* = src/kernel.cu (CopyToOutput, and the includes of kernel.cuh, scan.cuh, sky.cuh, pathtrace.cuh) */
#include "temporalDenoising.cuh"
#include "postprocessing.cuh"
#include <cstring>

GlobalSettings* g_settings = nullptr;      /* extern in src/kernel.cu:8 (defined in src/main.cu:43, which is not part of this build) */
/* RayTracer::draw / UpdateFrame come along with src/kernel.cu; nothing here calls them.  The members they reference live in
 * reference .cu files that are not part of this build (bvh.cu, denoising.cu, postprocessing.cu, inputControl.cu) — their launch
 * sequences are reproduced by the extern "C" functions below — so the loader only needs the symbols to exist. */
void RayTracer::BuildBvhLevel1() {}
void RayTracer::BuildBvhLevel2() {}
void RayTracer::PostProcessing(Int2, Int2) {}
void RayTracer::InputControlUpdate() {}
void RayTracer::TemporalSpatialDenoising(Int2, Int2) {}

static_assert(sizeof(Camera) == 176 && sizeof(SurfaceMaterial) == 32 && sizeof(Triangle) == 128 && sizeof(BVHNode) == 64,
              "type shim changed a reference layout");

#define CKF(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "refcuda(frame): %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return -1; } } while (0)

namespace {

enum RefBuf { RB_COLOR, RB_ACCUM, RB_HISTORY, RB_NORMAL, RB_DEPTH, RB_HISTORY_DEPTH, RB_MOTION, RB_NOISE, RB_NOISE16, RB_SKY, RB_SUN, RB_ALBEDO,
              RB_COLOR4, RB_COLOR16, RB_COLOR64, RB_SCALED, RB_OUTPUT, RB_COUNT };

struct Surf { cudaArray* array = nullptr; SurfObj surf = 0; int w = 0, h = 0, bytes = 0; };

struct RefFrame {
    int W = 0, H = 0;                         // renderWidth/Height of the current frame (<= the surface size they were created with)
    int histW = 0, histH = 0;                 // historyRenderWidth/Height (src/kernel.cu:83-84)
    Surf buf[RB_COUNT];
    // textures: slot -> mip levels (src/texture.h:14-25)
    cudaArray* texArray[MipmapTextureCount][Mipmap::numMipLevels] = {};
    TextureAtlas hAtlas = {};
    TextureAtlas* dAtlas = nullptr;
    // scene
    SceneGeometry geom = {};
    SceneMaterial mat = {};
    SurfaceMaterial* dMaterials = nullptr; int* dMaterialsIdx = nullptr;
    BlueNoiseRandGenerator randGen;
    unsigned char* dTables[3] = {};
    float *skyPdf = nullptr, *skyCdf = nullptr, *skyTmp = nullptr, *sunPdf = nullptr, *sunCdf = nullptr, *sunTmp = nullptr;
    ConstBuffer cbo = {};
    bool skyConstInit = false;
    // post-processing (src/init.cu:331-336): exposure[4], histogram[64], the output surface handle in device memory
    float* dExposure = nullptr; uint* dHistogram = nullptr; SurfObj* dRenderTarget = nullptr; int outW = 0, outH = 0;
};

int make_surf(Surf& s, cudaChannelFormatDesc d, int w, int h, int bytes) {
    s.w = w; s.h = h; s.bytes = bytes;
    CKF(cudaMallocArray(&s.array, &d, w, h, cudaArraySurfaceLoadStore));
    cudaResourceDesc r = {}; r.resType = cudaResourceTypeArray; r.res.array.array = s.array;
    CKF(cudaCreateSurfaceObject(&s.surf, &r));
    std::vector<unsigned char> z((size_t)w * h * bytes, 0);
    CKF(cudaMemcpy2DToArray(s.array, 0, 0, z.data(), (size_t)w * bytes, (size_t)w * bytes, h, cudaMemcpyHostToDevice));
    return 0;
}

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

}  // namespace

extern "C" {

int refcuda_frame_create(int W, int H, void** out) {
    RefFrame* f = new RefFrame();
    f->W = W; f->H = H; f->histW = W; f->histH = H;
    const cudaChannelFormatDesc h4 = cudaCreateChannelDescHalf4(), h2 = cudaCreateChannelDescHalf2(), h1 = cudaCreateChannelDescHalf1(),
                                f4 = cudaCreateChannelDesc<float4>();
    // src/init.cu:473-500
    if (make_surf(f->buf[RB_COLOR], h4, W, H, 8) || make_surf(f->buf[RB_ACCUM], h4, W, H, 8) || make_surf(f->buf[RB_HISTORY], h4, W, H, 8) ||
        make_surf(f->buf[RB_NORMAL], h4, W, H, 8) || make_surf(f->buf[RB_ALBEDO], h4, W, H, 8) || make_surf(f->buf[RB_DEPTH], h1, W, H, 2) ||
        make_surf(f->buf[RB_HISTORY_DEPTH], h1, W, H, 2) || make_surf(f->buf[RB_MOTION], h2, W, H, 4) ||
        make_surf(f->buf[RB_NOISE], h1, ceil_div(W, 8), ceil_div(H, 8), 2) || make_surf(f->buf[RB_NOISE16], h1, ceil_div(W, 16), ceil_div(H, 16), 2) ||
        make_surf(f->buf[RB_SKY], f4, SKY_WIDTH, SKY_HEIGHT, 16) || make_surf(f->buf[RB_SUN], f4, SKY_WIDTH, SKY_HEIGHT, 16)) return -1;
    CKF(cudaMalloc(&f->skyPdf, SKY_SIZE * sizeof(float))); CKF(cudaMalloc(&f->skyCdf, SKY_SIZE * sizeof(float)));
    CKF(cudaMalloc(&f->skyTmp, SKY_SCAN_BLOCK_COUNT * sizeof(float)));
    CKF(cudaMalloc(&f->sunPdf, SUN_SIZE * sizeof(float))); CKF(cudaMalloc(&f->sunCdf, SUN_SIZE * sizeof(float)));
    CKF(cudaMalloc(&f->sunTmp, SUN_SCAN_BLOCK_COUNT * sizeof(float)));
    CKF(cudaMalloc(&f->dAtlas, sizeof(TextureAtlas)));
    f->cbo.bvhBatchSize = 1024;                               // src/init.cu:191
    f->cbo.sampleParams = SampleParams();
    f->cbo.clockTime = 0.0f;
    *out = f;
    return 0;
}

void refcuda_frame_destroy(void* h) {
    RefFrame* f = (RefFrame*)h;
    if (!f) return;
    for (auto& s : f->buf) { if (s.surf) cudaDestroySurfaceObject(s.surf); if (s.array) cudaFreeArray(s.array); }
    for (int t = 0; t < MipmapTextureCount; ++t)
        for (int l = 0; l < (int)Mipmap::numMipLevels; ++l)
            if (f->texArray[t][l]) { cudaDestroySurfaceObject(f->hAtlas.mipmapTex[t].mip[l]); cudaFreeArray(f->texArray[t][l]); }
    cudaFree(f->skyPdf); cudaFree(f->skyCdf); cudaFree(f->skyTmp); cudaFree(f->sunPdf); cudaFree(f->sunCdf); cudaFree(f->sunTmp);
    cudaFree(f->dAtlas); cudaFree(f->dMaterials); cudaFree(f->dMaterialsIdx); cudaFree(f->dExposure); cudaFree(f->dHistogram); cudaFree(f->dRenderTarget);
    for (auto p : f->dTables) cudaFree(p);
    delete f;
}

// blue-noise tables (src/blueNoiseRandGenData.h), host pointers: sobol 65536, scrambling 131072, ranking 131072 bytes
int refcuda_frame_set_tables(void* h, const unsigned char* sobol, const unsigned char* scramble, const unsigned char* ranking) {
    RefFrame* f = (RefFrame*)h;
    const size_t n[3] = {256 * 256, 128 * 128 * 8, 128 * 128 * 8};
    const unsigned char* src[3] = {sobol, scramble, ranking};
    for (int i = 0; i < 3; ++i) { CKF(cudaMalloc(&f->dTables[i], n[i])); CKF(cudaMemcpy(f->dTables[i], src[i], n[i], cudaMemcpyHostToDevice)); }
    f->randGen.sobol_256spp_256d = f->dTables[0]; f->randGen.scramblingTile = f->dTables[1]; f->randGen.rankingTile = f->dTables[2];
    return 0;
}

// device pointers produced by refcuda_build_bvh; sizes as src/init.cu:191-194
int refcuda_frame_set_scene(void* h, void* dTriangles, void* dNodes, void* dTlas, int triCountPadded, int batchCount) {
    RefFrame* f = (RefFrame*)h;
    f->geom.aabbs = nullptr; f->geom.numAabbs = 0;
    f->geom.triangles = (Triangle*)dTriangles; f->geom.bvhNodes = (BVHNode*)dNodes; f->geom.tlasBvhNodes = (BVHNode*)dTlas;
    f->geom.numTriangles = triCountPadded;
    f->cbo.bvhNodesSize = triCountPadded; f->cbo.trianglesSize = triCountPadded; f->cbo.tlasBvhNodesSize = batchCount;
    return 0;
}

int refcuda_frame_set_materials(void* h, const void* materials, int n, const int* idx, int nIdx) {
    RefFrame* f = (RefFrame*)h;
    cudaFree(f->dMaterials); cudaFree(f->dMaterialsIdx);
    CKF(cudaMalloc(&f->dMaterials, (size_t)n * sizeof(SurfaceMaterial))); CKF(cudaMemcpy(f->dMaterials, materials, (size_t)n * sizeof(SurfaceMaterial), cudaMemcpyHostToDevice));
    CKF(cudaMalloc(&f->dMaterialsIdx, (size_t)nIdx * sizeof(int))); CKF(cudaMemcpy(f->dMaterialsIdx, idx, (size_t)nIdx * sizeof(int), cudaMemcpyHostToDevice));
    f->mat.materials = f->dMaterials; f->mat.materialsIdx = f->dMaterialsIdx; f->mat.numMaterials = n; f->mat.numMaterialsIdx = nIdx;
    return 0;
}

// one mip level of texture slot `slot` (rgba16, host).  Levels not supplied alias the last supplied one; maxLod = 10 (src/init.cu:571).
int refcuda_frame_set_texture_level(void* h, int slot, int level, int w, int hgt, const unsigned short* rgba) {
    RefFrame* f = (RefFrame*)h;
    if (slot < 0 || slot >= MipmapTextureCount || level < 0 || level >= (int)Mipmap::numMipLevels) return -2;
    const cudaChannelFormatDesc d = cudaCreateChannelDesc<ushort4>();
    if (f->texArray[slot][level]) { cudaDestroySurfaceObject(f->hAtlas.mipmapTex[slot].mip[level]); cudaFreeArray(f->texArray[slot][level]); }
    CKF(cudaMallocArray(&f->texArray[slot][level], &d, w, hgt, cudaArraySurfaceLoadStore));
    CKF(cudaMemcpy2DToArray(f->texArray[slot][level], 0, 0, rgba, (size_t)w * 8, (size_t)w * 8, hgt, cudaMemcpyHostToDevice));
    cudaResourceDesc r = {}; r.resType = cudaResourceTypeArray; r.res.array.array = f->texArray[slot][level];
    CKF(cudaCreateSurfaceObject(&f->hAtlas.mipmapTex[slot].mip[level], &r));
    f->hAtlas.mipmapTex[slot].size[level] = Int2(w, hgt);
    for (int l = level + 1; l < (int)Mipmap::numMipLevels; ++l)
        if (!f->texArray[slot][l]) { f->hAtlas.mipmapTex[slot].mip[l] = f->hAtlas.mipmapTex[slot].mip[level]; f->hAtlas.mipmapTex[slot].size[l] = Int2(w, hgt); }
    f->hAtlas.mipmapTex[slot].maxLod = 10;
    CKF(cudaMemcpy(f->dAtlas, &f->hAtlas, sizeof(TextureAtlas), cudaMemcpyHostToDevice));
    return 0;
}

// Sky / SkySun / Scan exactly as RayTracer::draw does on a regenerate (src/kernel.cu:285-308)
int refcuda_frame_generate_sky(void* h, const float* sunDir3, float timeOfDay, float sunAxisAngle, float skyScalar, float sunScalar, float sunAngle) {
    RefFrame* f = (RefFrame*)h;
    SkyParams sp; sp.timeOfDay = timeOfDay; sp.sunAxisAngle = sunAxisAngle; sp.skyScalar = skyScalar; sp.sunScalar = sunScalar; sp.sunAngle = sunAngle;
    const Float3 sunDir(sunDir3[0], sunDir3[1], sunDir3[2]);
    if (!f->skyConstInit) { InitSkyConstantBuffer(); f->skyConstInit = true; }
    sp.sunScalar = max(sp.sunScalar, 0.00001f); sp.skyScalar = max(sp.skyScalar, 0.00001f); sp.sunAngle = max(sp.sunAngle, 0.51f);
    UpdateSkyState(sunDir);
    Sky<<<dim3(SKY_WIDTH / 8, SKY_HEIGHT / 8, 1), dim3(8, 8, 1)>>>(f->buf[RB_SKY].surf, f->skyPdf, Int2(SKY_WIDTH, SKY_HEIGHT), sunDir, sp);
    Scan(f->skyPdf, f->skyCdf, f->skyTmp, SKY_SIZE, SKY_SCAN_BLOCK_SIZE, 1);
    SkySun<<<dim3(SUN_WIDTH / 8, SUN_HEIGHT / 8, 1), dim3(8, 8, 1)>>>(f->buf[RB_SUN].surf, f->sunPdf, Int2(SUN_WIDTH, SUN_HEIGHT), sunDir, sp);
    Scan(f->sunPdf, f->sunCdf, f->sunTmp, SUN_SIZE, SUN_SCAN_BLOCK_SIZE, 1);
    const float sunRadiusRadian = sp.sunAngle * M_PI / 180.0f / 2.0f;
    f->cbo.sunArea = powf(tanf(sunRadiusRadian), 2.0f) * M_PI;
    f->cbo.sunAngleCosThetaMax = cosf(sunRadiusRadian);
    f->cbo.sunDir = sunDir;
    CKF(cudaDeviceSynchronize()); CKF(cudaGetLastError());
    return 0;
}

// alternative to generate_sky: take the environment inputs from outside (host arrays): sky float4[512*256], sun float4[32*32]
int refcuda_frame_set_sky(void* h, const float* sky4, const float* skyCdf, const float* sun4, const float* sunCdf, const float* sunDir3,
                          float cosThetaMax, float sunArea) {
    RefFrame* f = (RefFrame*)h;
    CKF(cudaMemcpy2DToArray(f->buf[RB_SKY].array, 0, 0, sky4, SKY_WIDTH * 16, SKY_WIDTH * 16, SKY_HEIGHT, cudaMemcpyHostToDevice));
    CKF(cudaMemcpy2DToArray(f->buf[RB_SUN].array, 0, 0, sun4, SUN_WIDTH * 16, SUN_WIDTH * 16, SUN_HEIGHT, cudaMemcpyHostToDevice));
    CKF(cudaMemcpy(f->skyCdf, skyCdf, SKY_SIZE * sizeof(float), cudaMemcpyHostToDevice));
    CKF(cudaMemcpy(f->sunCdf, sunCdf, SUN_SIZE * sizeof(float), cudaMemcpyHostToDevice));
    f->cbo.sunDir = Float3(sunDir3[0], sunDir3[1], sunDir3[2]); f->cbo.sunAngleCosThetaMax = cosThetaMax; f->cbo.sunArea = sunArea;
    return 0;
}

int refcuda_frame_get_sky_cdfs(void* h, float* skyCdf, float* sunCdf, float* cosThetaMax) {
    RefFrame* f = (RefFrame*)h;
    CKF(cudaMemcpy(skyCdf, f->skyCdf, SKY_SIZE * sizeof(float), cudaMemcpyDeviceToHost));
    CKF(cudaMemcpy(sunCdf, f->sunCdf, SUN_SIZE * sizeof(float), cudaMemcpyDeviceToHost));
    *cosThetaMax = f->cbo.sunAngleCosThetaMax;
    return 0;
}

// PathTrace launch of RayTracer::draw (src/kernel.cu:334-350).  camera / historyCamera: sizeof(Camera) blobs (the camera.bin format).
int refcuda_frame_pathtrace(void* h, const void* camera, const void* historyCamera, int frameNum, int iterations, float* ms_out) {
    RefFrame* f = (RefFrame*)h;
    Camera cam, hist;
    memcpy(&cam, camera, sizeof(Camera)); memcpy(&hist, historyCamera ? historyCamera : camera, sizeof(Camera));
    f->cbo.camera = cam;
    f->cbo.historyCamera.Setup(hist);                                   // src/kernel.cu:135,357
    f->cbo.frameNum = frameNum;
    const int W = f->W, H = f->H;
    cudaEvent_t e0, e1; CKF(cudaEventCreate(&e0)); CKF(cudaEventCreate(&e1));
    float total = 0.0f;
    for (int it = 0; it < iterations; ++it) {
        CKF(cudaDeviceSynchronize());
        CKF(cudaEventRecord(e0));
        PathTrace<<<dim3(divRoundUp(W, 8), divRoundUp(H, 8), 1), dim3(8, 8, 1)>>>(
            f->cbo, f->geom, f->mat, f->randGen, f->buf[RB_COLOR].surf, f->buf[RB_NORMAL].surf, f->buf[RB_DEPTH].surf, f->buf[RB_SKY].surf, f->skyCdf,
            f->buf[RB_SUN].surf, f->sunCdf, f->buf[RB_MOTION].surf, f->buf[RB_NOISE].surf, f->buf[RB_ALBEDO].surf, f->dAtlas, Int2(W, H));
        CKF(cudaEventRecord(e1));
        CKF(cudaEventSynchronize(e1));
        CKF(cudaGetLastError());
        float ms = 0; CKF(cudaEventElapsedTime(&ms, e0, e1));
        if (it > 0 || iterations == 1) total += ms;
    }
    if (ms_out) ms_out[0] = total / (iterations > 1 ? iterations - 1 : 1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return 0;
}

// TemporalSpatialDenoising (src/denoising.cu:48-186): same kernels, same order, same launch shapes, no per-stage sync.
// passes = {enableTemporalDenoising, enableLocalSpatialFilter, enableWideSpatialFilter, enableTemporalDenoising2}; params = the 11 floats
// of DenoisingParams in declaration order (src/settingParams.h:144-157) or NULL for the defaults.
int refcuda_frame_denoise(void* h, int frameNum, const int* passes, const float* params, float* ms_out) {
    RefFrame* f = (RefFrame*)h;
    const int W = f->W, H = f->H;
    const Int2 bufferDim(W, H), historyDim(f->histW, f->histH);
    DenoisingParams dp;
    if (params) memcpy(&dp.local_denoise_sigma_normal, params, 11 * sizeof(float));
    f->cbo.frameNum = frameNum;
    const ConstBuffer& cbo = f->cbo;
    auto S = [&](int b) { return f->buf[b].surf; };
    const dim3 blockDim8(8, 8, 1), gridDim8(divRoundUp(W, 8), divRoundUp(H, 8), 1), grid16(divRoundUp(W, 16), divRoundUp(H, 16), 1), block16(16, 16, 1);
    const UInt2 n16(divRoundUp(W, 16), divRoundUp(H, 16));
    cudaEvent_t e0, e1; CKF(cudaEventCreate(&e0)); CKF(cudaEventCreate(&e1));
    CKF(cudaDeviceSynchronize());
    CKF(cudaEventRecord(e0));
    if (passes[0] && frameNum != 1)
        TemporalFilter<<<gridDim8, blockDim8>>>(cbo, S(RB_COLOR), S(RB_ACCUM), S(RB_NORMAL), S(RB_DEPTH), S(RB_HISTORY_DEPTH), S(RB_MOTION), dp, bufferDim, historyDim);
    if (passes[1]) {
        CalculateTileNoiseLevel<<<gridDim8, dim3(8, 4, 1)>>>(S(RB_COLOR), S(RB_DEPTH), S(RB_NOISE), bufferDim);
        TileNoiseLevel8x8to16x16<<<dim3(divRoundUp(n16.x, 8), divRoundUp(n16.y, 8), 1), blockDim8>>>(S(RB_NOISE), S(RB_NOISE16));
        SpatialFilter7x7<<<grid16, block16>>>(cbo, S(RB_COLOR), S(RB_NORMAL), S(RB_DEPTH), S(RB_NOISE16), dp, bufferDim);
    }
    if (passes[0]) CopyToHistoryColorBuffer<<<gridDim8, blockDim8>>>(S(RB_COLOR), S(RB_ACCUM), bufferDim);
    if (passes[2]) {
        CalculateTileNoiseLevel<<<gridDim8, dim3(8, 4, 1)>>>(S(RB_COLOR), S(RB_DEPTH), S(RB_NOISE), bufferDim);
        TileNoiseLevel8x8to16x16<<<dim3(divRoundUp(n16.x, 8), divRoundUp(n16.y, 8), 1), blockDim8>>>(S(RB_NOISE), S(RB_NOISE16));
        if (passes[2] == 2) {      // BASELINE config 4's 5-iteration a-trous schedule: the reference's kernel template with strides 1, 2, 4, 8, 16
            SpatialFilterGlobal5x5<1><<<grid16, block16>>>(cbo, S(RB_COLOR), S(RB_NORMAL), S(RB_DEPTH), S(RB_NOISE16), dp, bufferDim);
            SpatialFilterGlobal5x5<2><<<grid16, block16>>>(cbo, S(RB_COLOR), S(RB_NORMAL), S(RB_DEPTH), S(RB_NOISE16), dp, bufferDim);
            SpatialFilterGlobal5x5<4><<<grid16, block16>>>(cbo, S(RB_COLOR), S(RB_NORMAL), S(RB_DEPTH), S(RB_NOISE16), dp, bufferDim);
            SpatialFilterGlobal5x5<8><<<grid16, block16>>>(cbo, S(RB_COLOR), S(RB_NORMAL), S(RB_DEPTH), S(RB_NOISE16), dp, bufferDim);
            SpatialFilterGlobal5x5<16><<<grid16, block16>>>(cbo, S(RB_COLOR), S(RB_NORMAL), S(RB_DEPTH), S(RB_NOISE16), dp, bufferDim);
        } else {
        SpatialFilterGlobal5x5<3><<<grid16, block16>>>(cbo, S(RB_COLOR), S(RB_NORMAL), S(RB_DEPTH), S(RB_NOISE16), dp, bufferDim);
        SpatialFilterGlobal5x5<6><<<grid16, block16>>>(cbo, S(RB_COLOR), S(RB_NORMAL), S(RB_DEPTH), S(RB_NOISE16), dp, bufferDim);
        SpatialFilterGlobal5x5<12><<<grid16, block16>>>(cbo, S(RB_COLOR), S(RB_NORMAL), S(RB_DEPTH), S(RB_NOISE16), dp, bufferDim);
        }
    }
    ApplyAlbedo<<<gridDim8, blockDim8>>>(S(RB_COLOR), S(RB_ALBEDO), bufferDim);
    if (passes[3]) {
        if (frameNum != 1)
            TemporalFilter2<<<gridDim8, blockDim8>>>(cbo, S(RB_COLOR), S(RB_HISTORY), S(RB_NORMAL), S(RB_DEPTH), S(RB_HISTORY_DEPTH), S(RB_MOTION), S(RB_NOISE), bufferDim, historyDim);
        CopyToHistoryColorDepthBuffer<<<gridDim8, blockDim8>>>(S(RB_COLOR), S(RB_DEPTH), S(RB_HISTORY), S(RB_HISTORY_DEPTH), bufferDim);
    }
    f->histW = W; f->histH = H;               // historyRenderWidth/Height of the next frame (src/kernel.cu:83-84)
    CKF(cudaEventRecord(e1));
    CKF(cudaEventSynchronize(e1));
    CKF(cudaGetLastError());
    float ms = 0; CKF(cudaEventElapsedTime(&ms, e0, e1));
    if (ms_out) ms_out[0] = ms;
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return 0;
}

// RayTracer::PostProcessing with the default pass set (src/postprocessing.cu:16-158: DownScale4 x3, Histogram2, AutoExposure, BicubicScale,
// SharpeningFilter, ToneMappingReinhardExtended) + the CopyToOutput launch of draw() (src/kernel.cu:369-375), same launch shapes.
// params = {gain, maxWhite, gamma}; the exposure state persists across calls like d_exposure.
int refcuda_frame_postprocess(void* h, int frameNum, float deltaTime, int outW, int outH, const float* params, float* ms_out) {
    RefFrame* f = (RefFrame*)h;
    const int W = f->W, H = f->H;
    const UInt2 s4(divRoundUp(W, 4u), divRoundUp(H, 4u)), s16(divRoundUp(s4.x, 4u), divRoundUp(s4.y, 4u)), s64(divRoundUp(s16.x, 4u), divRoundUp(s16.y, 4u));
    if (!f->dExposure || f->outW != outW || f->outH != outH) {
        const cudaChannelFormatDesc h4 = cudaCreateChannelDescHalf4(), u4 = cudaCreateChannelDesc<uchar4>();
        for (int b : {RB_COLOR4, RB_COLOR16, RB_COLOR64, RB_SCALED, RB_OUTPUT})
            if (f->buf[b].surf) { cudaDestroySurfaceObject(f->buf[b].surf); cudaFreeArray(f->buf[b].array); f->buf[b] = Surf(); }
        if (make_surf(f->buf[RB_COLOR4], h4, s4.x, s4.y, 8) || make_surf(f->buf[RB_COLOR16], h4, s16.x, s16.y, 8) || make_surf(f->buf[RB_COLOR64], h4, s64.x, s64.y, 8) ||
            make_surf(f->buf[RB_SCALED], h4, outW, outH, 8) || make_surf(f->buf[RB_OUTPUT], u4, outW, outH, 4)) return -1;
        if (!f->dExposure) {
            CKF(cudaMalloc(&f->dExposure, 4 * sizeof(float))); CKF(cudaMalloc(&f->dHistogram, 64 * sizeof(uint))); CKF(cudaMalloc(&f->dRenderTarget, sizeof(SurfObj)));
            const float init[4] = {1.0f, 1.0f, 1.0f, 1.0f};                                            // src/init.cu:332-333
            CKF(cudaMemcpy(f->dExposure, init, sizeof(init), cudaMemcpyHostToDevice));
        }
        CKF(cudaMemcpy(f->dRenderTarget, &f->buf[RB_OUTPUT].surf, sizeof(SurfObj), cudaMemcpyHostToDevice));
        f->outW = outW; f->outH = outH;
    }
    PostProcessParams pp; pp.gain = params[0]; pp.maxWhite = params[1]; pp.gamma = params[2];
    f->cbo.frameNum = frameNum;
    const Int2 bufferDim(W, H), outputDim(outW, outH);
    const dim3 blockDim8(8, 8, 1), gridDim8(divRoundUp(W, 8), divRoundUp(H, 8), 1), gridDim4(divRoundUp(s4.x, 8u), divRoundUp(s4.y, 8u), 1),
               gridDim16(divRoundUp(s16.x, 8u), divRoundUp(s16.y, 8u), 1), scaleGrid(divRoundUp(outW, 8), divRoundUp(outH, 8), 1);
    auto S = [&](int b) { return f->buf[b].surf; };
    cudaEvent_t e0, e1; CKF(cudaEventCreate(&e0)); CKF(cudaEventCreate(&e1));
    CKF(cudaMemset(f->dHistogram, 0, 64 * sizeof(uint)));                                             // src/kernel.cu:278
    CKF(cudaDeviceSynchronize());
    CKF(cudaEventRecord(e0));
    DownScale4<<<gridDim8, blockDim8>>>(S(RB_COLOR), S(RB_COLOR4), bufferDim);
    DownScale4<<<gridDim4, blockDim8>>>(S(RB_COLOR4), S(RB_COLOR16), Int2(s4.x, s4.y));
    DownScale4<<<gridDim16, blockDim8>>>(S(RB_COLOR16), S(RB_COLOR64), Int2(s16.x, s16.y));
    Histogram2<<<1, dim3(min(s64.x, 32u), min(s64.y, 32u), 1)>>>(f->dHistogram, S(RB_COLOR64), Int2(s64.x, s64.y));
    AutoExposure<<<1, 1>>>(f->dExposure, f->dHistogram, (float)(s64.x * s64.y), deltaTime, pp.gain);
    BicubicScale<<<scaleGrid, blockDim8>>>(S(RB_SCALED), S(RB_COLOR), outputDim, bufferDim);
    SharpeningFilter<<<scaleGrid, blockDim8>>>(S(RB_SCALED), outputDim);
    ToneMappingReinhardExtended<<<scaleGrid, blockDim8>>>(S(RB_SCALED), outputDim, f->dExposure, pp);
    CopyToOutput<<<scaleGrid, blockDim8>>>(f->dRenderTarget, S(RB_SCALED), f->randGen, f->cbo, outputDim);
    CKF(cudaEventRecord(e1));
    CKF(cudaEventSynchronize(e1));
    CKF(cudaGetLastError());
    float ms = 0; CKF(cudaEventElapsedTime(&ms, e0, e1));
    if (ms_out) ms_out[0] = ms;
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return 0;
}

int refcuda_frame_get_exposure(void* h, float* exposure4, unsigned int* histogram64) {
    RefFrame* f = (RefFrame*)h;
    if (!f->dExposure) return -2;
    CKF(cudaMemcpy(exposure4, f->dExposure, 4 * sizeof(float), cudaMemcpyDeviceToHost));
    CKF(cudaMemcpy(histogram64, f->dHistogram, 64 * sizeof(uint), cudaMemcpyDeviceToHost));
    return 0;
}

// Dynamic resolution: UpdateFrame (src/kernel.cu:83-101) keeps the surfaces at their creation size and renders the next frame into
// their top-left renderWidth x renderHeight corner; the previous size becomes historyDim for TemporalSpatialDenoising (src/kernel.cu:266,360).
int refcuda_frame_set_render_size(void* h, int W, int H) {
    RefFrame* f = (RefFrame*)h;
    if (W < 1 || H < 1 || W > f->buf[RB_COLOR].w || H > f->buf[RB_COLOR].h) return -2;
    f->W = W; f->H = H;                      // historyDim stays the size of the last denoised frame (set at the end of refcuda_frame_denoise)
    return 0;
}

// top-left w x hgt corner of a surface <-> compact host rows
int refcuda_frame_read_region(void* h, int which, void* host, int w, int hgt) {
    RefFrame* f = (RefFrame*)h;
    if (which < 0 || which >= RB_COUNT) return -2;
    const Surf& s = f->buf[which];
    if (w < 1 || hgt < 1 || w > s.w || hgt > s.h) return -3;
    CKF(cudaDeviceSynchronize());
    CKF(cudaMemcpy2DFromArray(host, (size_t)w * s.bytes, s.array, 0, 0, (size_t)w * s.bytes, hgt, cudaMemcpyDeviceToHost));
    return 0;
}

int refcuda_frame_write_region(void* h, int which, const void* host, int w, int hgt) {
    RefFrame* f = (RefFrame*)h;
    if (which < 0 || which >= RB_COUNT) return -2;
    const Surf& s = f->buf[which];
    if (w < 1 || hgt < 1 || w > s.w || hgt > s.h) return -3;
    CKF(cudaMemcpy2DToArray(s.array, 0, 0, host, (size_t)w * s.bytes, (size_t)w * s.bytes, hgt, cudaMemcpyHostToDevice));
    return 0;
}

// surface <-> host (row-major, no pitch).  which = RefBuf.
int refcuda_frame_read(void* h, int which, void* host, size_t bytes) {
    RefFrame* f = (RefFrame*)h;
    if (which < 0 || which >= RB_COUNT) return -2;
    const Surf& s = f->buf[which];
    int hh = s.h;
    if (which == RB_SUN) hh = SUN_HEIGHT;
    const int ww = which == RB_SUN ? SUN_WIDTH : s.w;
    if (bytes != (size_t)ww * hh * s.bytes) return -3;
    CKF(cudaDeviceSynchronize());
    CKF(cudaMemcpy2DFromArray(host, (size_t)ww * s.bytes, s.array, 0, 0, (size_t)ww * s.bytes, hh, cudaMemcpyDeviceToHost));
    return 0;
}

int refcuda_frame_write(void* h, int which, const void* host, size_t bytes) {
    RefFrame* f = (RefFrame*)h;
    if (which < 0 || which >= RB_COUNT) return -2;
    const Surf& s = f->buf[which];
    if (bytes != (size_t)s.w * s.h * s.bytes) return -3;
    CKF(cudaMemcpy2DToArray(s.array, 0, 0, host, (size_t)s.w * s.bytes, (size_t)s.w * s.bytes, s.h, cudaMemcpyHostToDevice));
    return 0;
}

}  // extern "C"
