/* rtrt.h — C ABI of the B200-native hot path (LBVH build -> BVH traversal / path trace -> denoise).
 *
 * This is the drop-in boundary for the per-frame GPU hot path of wangkepfe/Real-Time-Ray-Tracing.
 * The reference has no FFI layer: its boundary is the C++ class `RayTracer` (src/kernel.cuh:431-621)
 * and the kernels it launches.  Every entry point below names the reference interface it replaces
 * (file:line relative to the reference root).  Plain pointers and sizes only; no C++/torch types.
 *
 * Conventions
 *   - every function returns 0 (RTRT_OK) or a negative rtrt_status; nothing exits or aborts
 *     (the reference's GpuErrorCheck prints and exit()s: src/cudaError.cuh:6-13);
 *   - one context per GPU; a context is not thread-safe, distinct contexts are independent;
 *   - all work is enqueued on the context's stream; only rtrt_sync, rtrt_get_timings and the
 *     *_host entry points synchronise with the host;
 *   - device pointers returned by rtrt_get_buffer are borrowed, valid until the next rtrt_set_mesh*
 *     or rtrt_destroy.
 */
#ifndef RTRT_H
#define RTRT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTRT_ABI_VERSION 1

typedef enum rtrt_status {
    RTRT_OK = 0,
    RTRT_ERR_ARG = -1,       /* bad argument */
    RTRT_ERR_CUDA = -2,      /* a CUDA call failed; see rtrt_last_error */
    RTRT_ERR_CAPACITY = -3,  /* size outside supported range (reference: assert, src/init.cu:100-101,126) */
    RTRT_ERR_STATE = -4,     /* call order (e.g. traverse before build) */
    RTRT_ERR_IO = -5,        /* file / parse error */
    RTRT_ERR_NCCL = -6
} rtrt_status;

typedef struct rtrt_ctx rtrt_ctx;

/* GlobalSettings (src/globalSettings.h:5-22) as loaded by LoadConfig (src/configLoader.cpp:5-27).
 * Same keys, same defaults; `[optimziation]` is the (misspelt) table name the reference reads. */
typedef struct rtrt_config {
    int32_t width, height;                 /* [resolution] width,height       default 1920x1080 */
    char    inputMeshFileName[256];        /* [file]                                             */
    char    inputTextureFileNames[4][256]; /* [file] (first 4 entries kept)                      */
    int32_t numInputTextureFileNames;
    int32_t loadCameraAtInit;              /* [file] default false                               */
    char    inputCameraFileName[256];
    char    cameraSaveFileName[256];
    int32_t useDynamicResolution;          /* [optimziation] default true                        */
    float   targetFps;                     /* default 60                                         */
    int32_t maxWidth, maxHeight;           /* default 3840x2160                                  */
    int32_t minWidth, minHeight;           /* default 640x480                                    */
} rtrt_config;

/* Ray (src/geometry.h:40-50) and the observable result of one TraverseBvh call (src/traverse.h:107-125). */
typedef struct rtrt_ray { float orig[3]; float pad16; float dir[3]; float pad32; } rtrt_ray;
typedef struct rtrt_hit {
    float   t;    /* RayMax = 1e11 on miss (src/kernel.cuh:69)                                   */
    float   u, v; /* barycentrics as left by RayTriangleWatertight (src/geometry.cuh:406-472)    */
    int32_t tri;  /* objectIdx = batch*1024 + batch-local triangle id; -1 on miss                */
} rtrt_hit;

/* Per-stage device times of the most recent call of each stage, CUDA events on the context stream
 * (replaces Timer/ScopeTimer, src/timer.h:7-99). */
typedef struct rtrt_timings {
    float build_bvh_ms, traverse_ms, pathtrace_ms, denoise_ms;
    uint64_t kernel_launches;     /* kernels launched by this library since rtrt_create */
    uint64_t rays_traced;         /* TraverseBvh calls executed by the last pathtrace / traverse */
    uint64_t node_visits;         /* inner-node visits / leaf (triangle) tests of the last pathtrace; only counted  */
    uint64_t leaf_tests;          /* when option "count_visits" = 1 (adds register pressure, off by default)        */
    uint64_t rays_traced_total;   /* TraverseBvh calls of all rtrt_pathtrace calls since rtrt_create (device-side sum)  */
    /* option "kernel_timing" = 1: CUDA-event times of the kernels of the last rtrt_pathtrace (wavefront path):
     * [0] k_pt_primary  [1] k_pt_logic_first  [2] k_pt_trace round 1  [3] k_pt_logic round 1  [4] k_pt_trace round 2  [5] rest */
    float pt_kernel_ms[6];
    /* option "count_visits" = 1: node visits / leaf tests of the bounce rounds 1..4 of the last rtrt_pathtrace */
    uint32_t round_node_visits[4], round_leaf_tests[4], round_rays[4];
    float wide_build_ms;          /* option wide_bvh: the collapse to the wide layout inside the last rtrt_build_bvh (build_bvh_ms then covers the LBVH build only;
                                     0 when the collapse ran on the side stream, wide_bvh = 2) */
} rtrt_timings;

/* Buffers observable through rtrt_get_buffer.  The first block are the DebugPrintFile observation
 * points of src/bvh.cu:15-96; the second block are the Buffer2DName surfaces on the path
 * (src/kernel.cuh:286-330, src/init.cu:473-500) exposed as linear device buffers. */
typedef enum rtrt_buffer {
    RTRT_BUF_TRIANGLES = 0,      /* Triangle[B*1024]   128 B each  (triangles.csv)            */
    RTRT_BUF_AABBS,              /* AABB[B*1024]        24 B       (aabbs.csv)                */
    RTRT_BUF_MORTON_PRESORT,     /* uint32[B*1024]  (morton.csv)  only with keep_presort=1    */
    RTRT_BUF_MORTON,             /* uint32[B*1024]  sorted        (morton2.csv)               */
    RTRT_BUF_REORDER_IDX,        /* uint32[B*1024]                (reorderIdx.csv)            */
    RTRT_BUF_BVH_NODES,          /* BVHNode[B*1024]     64 B      (bvhNodes.csv)              */
    RTRT_BUF_TLAS_AABBS,         /* AABB[Bp]                      (TLAS_aabbs.csv)            */
    RTRT_BUF_TLAS_MORTON_PRESORT,/* uint32[Bp]                    (TLAS_morton.csv)           */
    RTRT_BUF_TLAS_MORTON,        /* uint32[Bp]                    (TLAS_morton2.csv)          */
    RTRT_BUF_TLAS_REORDER_IDX,   /* uint32[Bp]                    (TLAS_reorderIdx.csv)       */
    RTRT_BUF_TLAS_BVH_NODES,     /* BVHNode[Bp]                   (TLAS_bvhNodes.csv)         */
    RTRT_BUF_TRI_PERMUTATION,    /* uint32[N]: sorted position -> input triangle id; only with option global_sort=1     */
    RTRT_BUF_COUNT_STAGE1,
    /* Buffer2DName surfaces of the path (src/kernel.cuh:286-330), linear, row-major, W x H unless noted */
    RTRT_BUF_RENDER_COLOR = 32,  /* half3 colour (demodulated radiance) + uint16 material mask, 8 B/px  */
    RTRT_BUF_NORMAL,             /* half4, 8 B/px                                                       */
    RTRT_BUF_ALBEDO,             /* half4, 8 B/px                                                       */
    RTRT_BUF_DEPTH,              /* half, 2 B/px (RayMax stores as +inf)                                */
    RTRT_BUF_MOTION_VECTOR,      /* half2, 4 B/px (history uv - uv + 0.5)                               */
    RTRT_BUF_ACCUMULATION_COLOR, /* half3+u16, history of the spatially filtered colour                 */
    RTRT_BUF_HISTORY_COLOR,      /* half3+u16, history of the final colour                              */
    RTRT_BUF_HISTORY_DEPTH,      /* half                                                                */
    RTRT_BUF_NOISE_LEVEL,        /* half, ceil(W/8) x ceil(H/8)                                         */
    RTRT_BUF_NOISE_LEVEL_16,     /* half, ceil(W/16) x ceil(H/16)                                       */
    RTRT_BUF_SKY,                /* float4[512*256]  (SkyBuffer)                                        */
    RTRT_BUF_SKY_CDF,            /* float[512*256]   inclusive luminance CDF                            */
    RTRT_BUF_SUN,                /* float4[32*32]    (SunBuffer)                                        */
    RTRT_BUF_SUN_CDF,            /* float[32*32]                                                        */
    RTRT_BUF_ACCUM_RADIANCE,     /* float4: sum of colour*albedo over the accumulated samples, .w = count (rtrt_accumulate) */
    RTRT_BUF_OUTPUT_RGBA8,       /* uchar4[outW*outH]: the tone-mapped, dithered frame CopyToOutput writes (src/kernel.cu:26-59); rtrt_postprocess */
    RTRT_BUF_COLOR64,            /* half4[w64*h64]: ColorBuffer64, the 1/64-resolution colour the histogram reads                    */
    RTRT_BUF_EXPOSURE,           /* float[4]: d_exposure = EV, adapted luminance, adapted bright luminance, bright luminance          */
    RTRT_BUF_HISTOGRAM,          /* uint32[64]: d_histogram                                                                           */
    RTRT_BUF_WF_COUNTERS = 100,  /* diagnostics: uint32[64] of the wavefront path tracer: [r] rays of bounce round r, [16+r] hits, [24+r] misses,
                                    with option count_visits: [40+r] max and [48+r] sum of traversal iterations of round r */
    /* option wide_bvh = 1 (layout documented in csrc/wide_bvh.cuh): */
    RTRT_BUF_WIDE_NODES = 101,   /* 128 B per node slot: batch b owns slots [512 b, 512 b + 512), the top level follows at 512 B */
    RTRT_BUF_WIDE_TRIS = 102,    /* 48 B per triangle {v1.xyz, original index; v2.xyz, 0; v3.xyz, 0}, 1024 per batch, leaf-allocation order */
    RTRT_BUF_WIDE_STAT = 103     /* uint32[4]: error flags, nodes written, top-level nodes, 0 */
} rtrt_buffer;

/* SkyParams (src/settingParams.h:26-46), same defaults. */
typedef struct rtrt_sky_params { float timeOfDay, sunAxisAngle, skyScalar, sunScalar, sunAngle; } rtrt_sky_params;
/* DenoisingParams (src/settingParams.h:118-157), same defaults. */
typedef struct rtrt_denoise_params {
    float local_denoise_sigma_normal, local_denoise_sigma_depth, local_denoise_sigma_material;
    float large_denoise_sigma_normal, large_denoise_sigma_depth, large_denoise_sigma_material;
    float temporal_denoise_sigma_normal, temporal_denoise_sigma_depth, temporal_denoise_sigma_material;
    float noise_threshold_local, noise_threshold_large;
} rtrt_denoise_params;
/* the denoiser toggles of RenderPassSettings (src/settingParams.h:66-99).  enableWideSpatialFilter = 2 selects the 5-iteration a-trous
 * schedule of BASELINE config 4: five SpatialFilterGlobal5x5 passes with strides 1, 2, 4, 8, 16 (the kernel template of
 * src/temporalDenoising.cuh:495-608 with the classic power-of-two strides) instead of the shipped three with strides 3, 6, 12. */
typedef struct rtrt_pass_settings {
    int32_t enableTemporalDenoising, enableLocalSpatialFilter, enableWideSpatialFilter, enableTemporalDenoising2;
} rtrt_pass_settings;
/* PostProcessParams (src/settingParams.h:101-116: exposure 1, gain 40, maxWhite 7, gamma 2.2) and the post-processing toggles of
 * RenderPassSettings (src/settingParams.h:92-99).  Only the default tone mapper (Reinhard extended, luminance) is implemented;
 * bloom and lens flare (off by default in the reference) are not. */
typedef struct rtrt_postprocess_params {
    float exposure, gain, maxWhite, gamma;
    int32_t enablePostProcess, enableAutoExposure, enableSharpening, enableToneMapping;
} rtrt_postprocess_params;
/* SurfaceMaterial (src/kernel.cuh:182-196), 32 bytes; type = SurfaceMaterialType (src/kernel.cuh:172-180) */
typedef struct rtrt_material { float albedo[3]; uint32_t type; float F0[3]; float alpha; } rtrt_material;

/* ---- lifetime: RayTracer::RayTracer / init / cleanup (src/kernel.cuh:435-449, src/init.cu:53, :601-663) ---- */
int rtrt_create(const rtrt_config* cfg_or_null, int device, rtrt_ctx** out);
int rtrt_destroy(rtrt_ctx* ctx);
const char* rtrt_last_error(const rtrt_ctx* ctx);
int rtrt_abi_version(void);

/* ---- config: LoadConfig (src/configLoader.cpp:5-27); path is a parameter instead of the fixed
 *      "resources/config.toml".  Missing keys take the reference defaults. ---- */
int rtrt_config_defaults(rtrt_config* out);
int rtrt_load_config_toml(const char* path, rtrt_config* out);

/* ---- mesh: the upload block of RayTracer::init (src/init.cu:83-150).  Host arrays are copied, never
 *      retained.  pos/nrm: V x 3 floats (Float3), idx: N x 3 uint32.  nrm may be NULL: smooth normals are
 *      then generated on the device with GenerateSmoothNormals semantics (src/kernel.cu:228-257).
 *      Capacity: 1 <= N; the reference's caps (N <= 1M, < 1024 batches: src/kernel.cuh:54,
 *      src/init.cu:126) are lifted — see DESIGN.md "generalisation rule". ---- */
int rtrt_set_mesh(rtrt_ctx* ctx, const float* pos, const float* nrm_or_null, const uint32_t* idx,
                  uint32_t num_vertices, uint32_t num_triangles);
/* Same, with inputs already resident in device memory (dynamic geometry: the caller animates the
 * vertex buffer in place every frame, cf. MeshDisplace src/kernel.cu:139-226).  Pointers are borrowed. */
int rtrt_set_mesh_device(rtrt_ctx* ctx, const float* d_pos, const float* d_nrm, const uint32_t* d_idx,
                         uint32_t num_vertices, uint32_t num_triangles);
/* Mesh file `uint32 triCount; Triangle[triCount]` (LoadTrianglesFromFile src/init.cu:28-50, written by
 * tool/meshProcessor.cpp:204-208): 128-byte triangles, or the legacy 64-byte layout (positions only, e.g. the shipped
 * resources/models/test.bin; vertex normals are then generated) — told apart by the file size. */
int rtrt_set_mesh_bin(rtrt_ctx* ctx, const char* path);
/* Wavefront OBJ (the reference reads it through Assimp: LoadScene src/fileUtils.cu:16-80): v / vn / f, corners v, v/vt, v//vn,
 * v/vt/vn, negative indices, polygons fanned from the first corner; triangles in file order; normals used only when every
 * corner has one (mesh->HasNormals()), else generated.  RTRT_ERR_IO with the line number on malformed input. */
int rtrt_set_mesh_obj(rtrt_ctx* ctx, const char* path);
/* The offline mesh tool (tool/meshProcessor.cpp:146-215): writes the current mesh as a `.bin`.  presort != 0 orders the triangles by
 * the 60-bit Morton code of their centroid (computed and sorted on the device, stable); legacy64 != 0 writes the 64-byte layout. */
int rtrt_write_mesh_bin(rtrt_ctx* ctx, const char* path, int presort, int legacy64);
/* The two readers without a context (pure host code, usable without a GPU): `.obj` by extension, anything else `.bin`.
 * On success *tri_pos (and *tri_nrm, NULL when the file carries no normals) are malloc'ed arrays of 9 floats per triangle,
 * released with rtrt_mesh_file_free.  On failure the message is copied to err (may be NULL). */
int rtrt_mesh_file_read(const char* path, float** tri_pos, float** tri_nrm, uint32_t* num_tris, char* err, size_t err_len);
void rtrt_mesh_file_free(void* p);
/* Per-frame vertex animation from HOST arrays: replaces the contents of the vertex (and, if non-NULL, normal)
 * buffer of the current mesh, V x 3 floats each, asynchronously on the context stream (pinned memory makes the
 * copy overlap).  Topology is unchanged; call rtrt_build_bvh / rtrt_render_frame afterwards.  This is the
 * host-fed equivalent of MeshDisplace writing vertexBuffer every frame (src/kernel.cu:139-226, :324). */
int rtrt_update_vertices(rtrt_ctx* ctx, const float* pos, const float* nrm_or_null, uint32_t num_vertices);

/* ---- options ---- */
int rtrt_set_option_i(rtrt_ctx* ctx, const char* name, int64_t value);
/*   "keep_presort_morton" 0/1  : also fill RTRT_BUF_*MORTON_PRESORT (debug observation point)
 *   "wide_bvh"            0/1  : traversal layout.  0 = the reference's arrays (64-byte two-child nodes, one 128-byte triangle per leaf),
 *                                traversed with the reference's loop, caps and quirks: per-ray (tri, t, u, v) bit-identical to it.
 *                                1 = rtrt_build_bvh additionally derives an 8-wide BVH with fp16 child boxes relative to the node box (one 128-byte line
 *                                per visit) and 48-byte triangles from those arrays (they stay available and bit-exact), and rtrt_traverse /
 *                                rtrt_pathtrace use it: same watertight triangle test on the same vertex bits, conservative box test, spill
 *                                stack and no iteration cap -> (tri, t) bit-identical to the reference's search on every ray where the
 *                                reference neither drops a push on its full stack nor hits its iteration cap nor ties two triangles in t
 *   "stack_size"          16   : traversal stack entries (src/traverse.h:28): 1..16 for rtrt_pathtrace, 1..64 for rtrt_traverse
 *                                (> 32: stack in local memory; with a large max_loop this is the uncapped search used as a checker)
 *   "tlas_spin_limit"     4M   : polls before the top-level CTA of the fused build gives up waiting for the batch boxes; a time-out is
 *                                reported as RTRT_ERR_CUDA by the next rtrt_sync / rtrt_get_timings (0 forces it: test knob)
 *   "keep_all_triangles"  0/1  : 1 = the up-to-3 pad items of a partial last batch sort behind every real triangle, so no real triangle is
 *                                pushed out of the hierarchy (0 = the reference's behaviour, src/init.cu:103-115)
 *   "max_loop"            1024 : traversal iteration cap  (src/traverse.h:132)
 *   "pt_chunks"           1    : rtrt_pathtrace splits the frame into this many independent sets of 8-row strips and runs their
 *                                wavefront pipelines on side streams, so that the latency tail of one set's bounce rounds overlaps
 *                                the bulk of another's (1..8; results do not depend on it; measured slower than 1 on B200, kept for experiments)
 *   "any_hit_rays"        1    : inside rtrt_pathtrace (wavefront path), shadow rays and the rays of the last path segment stop at
 *                                the first accepted triangle: only "reaches the sky or not" is used of them (the reference completes
 *                                the closest-hit search and classifies afterwards, src/traverse.cuh:37-48).  Frame outputs are
 *                                identical with 0 and 1; 0 reproduces the reference's per-ray visit counts
 *   "assume_bvh_valid"    0/1  : mark the BVH arrays as built (they were filled from outside, e.g. an NCCL broadcast from
 *                                the rank that built them) or as stale
 *   "global_sort"         0/1  : rtrt_build_bvh first sorts the triangles by the 60-bit Morton code of their centroid in the
 *                                mesh bounding box (the offline step of tool/meshProcessor.cpp:146-200, on the device, per
 *                                call); triangle ids reported by the library are then positions in that order
 *   "morton_bits"         60   : code width of the global pre-sort: 60 = the reference tool's (20 bits per axis, two interleaved 30-bit
 *                                halves, tool/meshProcessor.cpp:52-62), 63 = 21 bits per axis in one interleave
 *   "wavefront"           1/0  : path-trace driver: wavefront ray queues with persistent traversal (default) or one megakernel
 *   "count_visits"        0/1  : rtrt_pathtrace also counts node visits / leaf tests (rtrt_timings)         */

/* ---- stage 1: RayTracer::BuildBvhLevel1 + BuildBvhLevel2 (src/bvh.cu:7-97) ---- */
int rtrt_build_bvh(rtrt_ctx* ctx);

/* ---- stage 2 (traversal only): TraverseBvh (src/traverse.h:107-253) with the state initialisation of
 *      RaySceneIntersect (src/traverse.cuh:86-101), one call per ray.  any_hit != 0 stops at the first
 *      accepted triangle (shadow-ray query; the reference classifies after a closest-hit,
 *      src/traverse.cuh:37-48).  d_rays / d_hits are DEVICE pointers. ---- */
int rtrt_traverse(rtrt_ctx* ctx, const rtrt_ray* d_rays, uint32_t n, rtrt_hit* d_hits, int any_hit);
/* Same through HOST buffers (H2D of rays, traversal, D2H of hits, synchronous). */
int rtrt_traverse_host(rtrt_ctx* ctx, const rtrt_ray* h_rays, uint32_t n, rtrt_hit* h_hits, int any_hit);
/* NO COUNTERPART IN THE REFERENCE — added for callers that cannot call cudaMalloc (CUDA graph
 * capture around their own work, tight serving loops, harnesses that forbid it).  rtrt_traverse_host without any allocation: the device
 * staging of rays / hits and the visit counters live in one file-scope `static __device__` array of csrc/traverse_noalloc.cu
 * (RTRT_NOALLOC_MAX_RAYS x 48 B + 256 B = 3 MiB, held by every GPU that loads the library until the process exits); its address is
 * taken with cudaGetSymbolAddress once per device.  n > RTRT_NOALLOC_MAX_RAYS -> RTRT_ERR_CAPACITY (split the batch).  Calls no
 * cudaMalloc / cudaFree: it always traverses the reference-layout arrays of the last rtrt_build_bvh (option wide_bvh is ignored, stack_size
 * <= 32), so the hits are bit-identical to rtrt_traverse_host with wide_bvh = 0.  *node_visits / *leaf_tests (may be NULL) receive the
 * inner-node visits / triangle tests of this call.  The array is shared by all calls and contexts on a GPU and keeps the last call's
 * contents (the counters are cleared per call); the call synchronises before returning, calls that overlap on one GPU (several host
 * threads) race. */
#define RTRT_NOALLOC_MAX_RAYS 65536u
int rtrt_traverse_host_noalloc(rtrt_ctx* ctx, const rtrt_ray* h_rays, uint32_t n, rtrt_hit* h_hits, int any_hit,
                               uint64_t* node_visits, uint64_t* leaf_tests);

/* ---- frame inputs -------------------------------------------------------------------------------
 * Camera: the raw `Camera` struct of src/kernel.cuh:78-101 (176 bytes), which is also the camera.bin file format
 * written by SaveCameraToFile (src/inputControl.cu:115-150); derived fields must be filled (Camera::update()).
 * The first rtrt_pathtrace after rtrt_set_camera(..., reset_history=1) uses the same camera as history camera
 * (src/kernel.cu:131-135); afterwards the history camera is the previous frame's (src/kernel.cu:357). */
int rtrt_set_camera(rtrt_ctx* ctx, const void* camera_blob, size_t bytes, int reset_history);
int rtrt_load_camera_file(rtrt_ctx* ctx, const char* path);          /* LoadCameraFromFile src/inputControl.cu:133-150 */
int rtrt_save_camera_file(rtrt_ctx* ctx, const char* path);          /* SaveCameraToFile   src/inputControl.cu:115-131 */
/* renderWidth/renderHeight of the next frame (src/kernel.cu:74,100).  Storage is grow-only: any size up to the largest one seen
 * (create the context at maxWidth x maxHeight like src/init.cu:55-61) re-uses the buffers and KEEPS the temporal history, which the
 * next rtrt_denoise samples with its own size as `historyDim` (src/kernel.cu:83-84,266; src/temporalDenoising.cuh:812,837).
 * Buffers read through rtrt_get_buffer are compact (row stride = the size they were written at). */
int rtrt_set_render_size(rtrt_ctx* ctx, int32_t width, int32_t height);
/* The dynamic-resolution controller of RayTracer::UpdateFrame (src/kernel.cu:78-101), as a pure function: given the current
 * render width and the last frame time it returns the next render size (width snapped to 16, clamped to [minWidth, maxWidth],
 * height = width/16*9).  With cfg->useDynamicResolution == 0 the size is returned unchanged (height from cur_height).
 * The caller applies it with rtrt_set_render_size and must set Camera.resolution to match (src/kernel.cu:100-101). */
int rtrt_dynamic_resolution_step(const rtrt_config* cfg, int32_t cur_width, int32_t cur_height, float delta_time_ms,
                                 int32_t* new_width, int32_t* new_height);
int rtrt_set_frame_number(rtrt_ctx* ctx, int32_t frame_num);         /* cbo.frameNum (seeds the blue-noise sample index) */
int rtrt_set_sky_params(rtrt_ctx* ctx, const rtrt_sky_params* p);    /* NULL = defaults; sky maps regenerate lazily (src/kernel.cu:289-308) */
int rtrt_get_sun(rtrt_ctx* ctx, float* sun_dir3, float* sun_angle_cos_theta_max);
int rtrt_set_denoise_params(rtrt_ctx* ctx, const rtrt_denoise_params* p, const rtrt_pass_settings* passes);
/* materials[num_materials] + per-triangle material index (NULL/0 = the reference scene: 10 materials, every
 * triangle -> material 3 = Lambert 0.9, src/init.cu:215-265) */
int rtrt_set_materials(rtrt_ctx* ctx, const rtrt_material* materials, int32_t num_materials,
                       const int32_t* materials_idx, uint32_t num_idx);
/* texture atlas slot 0 = albedo+AO, 1 = normal+roughness (src/texture.h:5-12): 16-bit RGBA, w x h, mip chain is
 * generated with the 2x2 box rule of MipmapGen (src/mipgen.cu:121-139).  Default: constant textures. */
int rtrt_set_texture(rtrt_ctx* ctx, int32_t slot, int32_t width, int32_t height, const uint16_t* rgba16);
/* rows [begin,end) of the frame this context traces and denoises (multi-GPU row bands); default: all rows */
int rtrt_set_row_range(rtrt_ctx* ctx, int32_t row_begin, int32_t row_end);
/* multi-GPU load balance: rtrt_pathtrace / rtrt_accumulate cover the rows with (row / strip_height) % num_shards == shard instead of
 * the contiguous band (a terrain frame is sky at the top, geometry at the bottom); rtrt_denoise keeps using the row range, whose
 * band + 80-row halo must have been made valid by the caller (all-gather).  num_shards <= 1 switches strips off. */
int rtrt_set_row_strips(rtrt_ctx* ctx, int32_t strip_height, int32_t num_shards, int32_t shard);
int rtrt_set_data_dir(rtrt_ctx* ctx, const char* dir);               /* where blue_noise_4spp.bin / hosek_sky.bin live */

/* ---- stage 2: the PathTrace launch of RayTracer::draw (src/kernel.cu:336-357, src/pathtrace.cuh:11-128) ---- */
int rtrt_pathtrace(rtrt_ctx* ctx);
/* ---- stage 3: RayTracer::TemporalSpatialDenoising (src/denoising.cu:5-188) ---- */
int rtrt_denoise(rtrt_ctx* ctx);
/* ---- progressive mode (BASELINE config 5, N spp): add the remodulated radiance (colour x albedo, the ApplyAlbedo step of
 *      src/temporalDenoising.cuh:1127-1139 in fp32) of the frame just path traced to RTRT_BUF_ACCUM_RADIANCE, rows of the
 *      current row range only; reset != 0 starts a new accumulation. ---- */
int rtrt_accumulate(rtrt_ctx* ctx, int reset);
/* ---- RayTracer::PostProcessing (src/postprocessing.cu:5-160, default passes) + CopyToOutput (src/kernel.cu:26-59,
 *      :369-375) on the frame in RTRT_BUF_RENDER_COLOR: down-scale x3, 64-bin histogram, auto exposure (delta_time_ms = the frame time
 *      the reference's Timer feeds the eye adaptation with), Catmull-Rom scale to the output size, sharpening, Reinhard-extended
 *      tone map + gamma, blue-noise dither to RGBA8 -> RTRT_BUF_OUTPUT_RGBA8.  Output size defaults to the render size. ---- */
int rtrt_set_postprocess_params(rtrt_ctx* ctx, const rtrt_postprocess_params* p_or_null_for_defaults);
int rtrt_set_output_size(rtrt_ctx* ctx, int32_t width, int32_t height);
int rtrt_postprocess(rtrt_ctx* ctx, float delta_time_ms);
/* binary PPM (P6) of RTRT_BUF_OUTPUT_RGBA8, like the reference's writeToPPM frame dump (src/kernel.cu:379-390); synchronises */
int rtrt_write_ppm(rtrt_ctx* ctx, const char* path);
/* ---- the three stages in draw() order, then frameNum++ (src/kernel.cu:259-362) ---- */
int rtrt_render_frame(rtrt_ctx* ctx);

/* ---- observation points: GetBuffer2D (src/kernel.cuh:455) + the DebugPrintFile dumps (src/bvh.cu) ---- */
int rtrt_get_buffer(rtrt_ctx* ctx, int which /* rtrt_buffer */, void** dptr, size_t* bytes);
/* convenience: synchronous copy of a buffer to host memory (dst must hold *bytes) */
int rtrt_read_buffer(rtrt_ctx* ctx, int which, void* dst, size_t dst_bytes);

int rtrt_get_counts(rtrt_ctx* ctx, uint32_t* num_triangles, uint32_t* num_batches, uint32_t* num_batches_padded);
/* Frames in flight.  rtrt_update_vertices and rtrt_read_buffer_async run their PCIe copies on a second (copy) stream, ordered
 * against the kernels by events: the upload waits for the last build to have consumed the vertex buffers and the next
 * rtrt_build_bvh waits for the upload; the read-back takes a device-side snapshot of the buffer on the main stream (after the
 * kernels that produce it) and copies the snapshot to `dst` on the copy stream while the next frame is already running.
 * `dst` (pinned memory for a real overlap) is valid after rtrt_sync, or after a stream/event wait that follows rtrt_join.
 * One read-back may be in flight per context; a second one waits for the first.  rtrt_join makes the main stream wait for all
 * queued copies, rtrt_sync blocks the host until both streams are idle.  (The reference presents through Vulkan interop and
 * never reads a frame back; this is the headless replacement of that hand-over, src/main.cu:1295-1395.) */
int rtrt_read_buffer_async(rtrt_ctx* ctx, int which, void* dst, size_t dst_bytes);
int rtrt_join(rtrt_ctx* ctx);
int rtrt_sync(rtrt_ctx* ctx);
int rtrt_get_timings(rtrt_ctx* ctx, rtrt_timings* out);   /* synchronises the stream */
/* ---- multi-GPU: SURVEY 8(b) `rtrt_create_multi`, 8(e) (csrc/multi.cu).  No counterpart in the reference (single-GPU); the stages are those of
 * RayTracer::draw (src/kernel.cu:259-362).  One context per GPU and rank (one process per GPU, or one host thread per context).  NCCL over
 * NVLink carries the vertex broadcast, the strip -> band redistribution of the G-buffers, the a-trous HALO EXCHANGE between the denoiser's
 * passes and the frame-buffer gather; NCCL is loaded with dlopen at the first use (RTRT_NCCL_LIB overrides the library path) and its errors
 * are reported as RTRT_ERR_NCCL.
 *   rtrt_multi_unique_id    ncclGetUniqueId: 128 bytes the host hands to every rank (MPI, torch.distributed, a file)
 *   rtrt_create_multi       rtrt_create + rtrt_multi_init
 *   rtrt_multi_init         ncclCommInitRank on the context's device; the context becomes rank `rank` of `world`
 *   rtrt_multi_attach_comm  the same with a caller-owned ncclComm_t (must come from the NCCL library this process uses)
 *   rtrt_multi_link_local   N contexts of ONE process become ranks 0..N-1 of an in-process group (copies between the contexts, host barriers;
 *                           each context must then be driven by its own host thread): single-process multi-GPU, and the single-GPU tests
 *   rtrt_multi_set_mode     RTRT_MULTI_ROWS (default): ONE frame, path-traced in interleaved strips of `strip_height` rows (strip k -> rank k % N)
 *                           and denoised in contiguous bands of whole 16-row tiles with a halo exchange per filter pass (strong scaling);
 *                           RTRT_MULTI_SAMPLES: every rank renders the whole frame with sample index frame + rank, the denoised frames are
 *                           averaged by one all-reduce (weak scaling: N samples per pixel and frame)
 *   rtrt_update_vertices_multi   rank 0 passes the host arrays (others NULL): one PCIe upload, NCCL broadcast of the vertex / normal buffers
 *   rtrt_render_frame_multi      one frame; on return RTRT_BUF_RENDER_COLOR holds the whole denoised frame on EVERY rank and the temporal history
 *                                is complete; frame number += 1 (rows) or += N (samples).  Collective: every rank must call it
 *   rtrt_render_progressive_multi  BASELINE config 5: spp samples accumulated in fp32, rank r renders samples r, r + N, ... of the whole frame, one
 *                                all-reduce (sum) of RTRT_BUF_ACCUM_RADIANCE at the end
 *   rtrt_multi_get_band     the rank's denoiser band, and the exchanges / bytes sent so far
 *   rtrt_multi_plan_rows    pure host function (no GPU): the rows `rank` sends to `peer` in exchange kind 0 (G-buffer redistribution: rows of
 *                           rank's strips inside peer's band +- 32), 1 (halo of `halo` rows around peer's band), 2 (frame gather); returns the
 *                           count and writes up to `cap` row numbers */
#define RTRT_NCCL_UNIQUE_ID_BYTES 128
#define RTRT_MULTI_ROWS 0
#define RTRT_MULTI_SAMPLES 1
int rtrt_multi_unique_id(void* id_out_128_bytes);
int rtrt_create_multi(const rtrt_config* cfg_or_null, int device, const void* unique_id, int rank, int world, rtrt_ctx** out);
int rtrt_multi_init(rtrt_ctx* ctx, const void* unique_id, int rank, int world);
int rtrt_multi_attach_comm(rtrt_ctx* ctx, void* nccl_comm, int rank, int world);
int rtrt_multi_link_local(rtrt_ctx** ctxs, int n);
int rtrt_multi_set_mode(rtrt_ctx* ctx, int mode, int strip_height);
int rtrt_update_vertices_multi(rtrt_ctx* ctx, const float* pos_rank0, const float* nrm_rank0_or_null, uint32_t num_vertices);
int rtrt_render_frame_multi(rtrt_ctx* ctx);
int rtrt_render_progressive_multi(rtrt_ctx* ctx, int32_t spp, int32_t first_frame);
int rtrt_multi_get_band(rtrt_ctx* ctx, int32_t* row_begin, int32_t* row_end, uint64_t* exchanges, uint64_t* bytes_sent);
int rtrt_multi_plan_rows(int32_t height, int32_t world, int32_t strip_height, int32_t rank, int32_t peer, int32_t kind, int32_t halo,
                         int32_t* rows_out, int32_t cap);

/* Measured ceilings of a traversal without RT cores on this GPU (csrc/microbench.cu; replaces nothing in the reference — it is the
 * denominator SURVEY 8(d) asks for: "node-fetch bytes/s vs measured L2 bandwidth"): L2 read bandwidth of all SMs streaming an
 * L2-resident 32 MiB buffer, and the latency of one dependent 128-byte-line load served by L1 (64 KiB set), L2 (16 MiB) and DRAM
 * (512 MiB).  Allocates and frees its own buffers, synchronises; not for timed regions. */
typedef struct rtrt_microbench_result {
    float l2_read_gbs;
    float l1_latency_ns, l2_latency_ns, dram_latency_ns;
    float l1_latency_cycles, l2_latency_cycles, dram_latency_cycles;
    float reserved_;
} rtrt_microbench_result;
int rtrt_microbench(rtrt_ctx* ctx, rtrt_microbench_result* out);
/* the cudaStream_t all work is enqueued on, as an opaque pointer (for event timing by the caller) */
void* rtrt_get_stream(rtrt_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* RTRT_H */
