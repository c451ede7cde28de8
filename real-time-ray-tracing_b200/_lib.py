"""ctypes binding of the C ABI declared in include/rtrt.h.

The library is the product: there is no Python/NumPy/CPU fallback.  If librtrt.so is missing or does
not load, importing this module raises — loudly, on purpose.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "librtrt.so")
if os.environ.get("RTRT_EXPERIMENT") == "1" and os.environ.get("RTRT_LIB"):   # A/B builds of the same library, experiments only
    LIB_PATH = os.environ["RTRT_LIB"]

RTRT_OK = 0
STATUS_NAMES = {0: "RTRT_OK", -1: "RTRT_ERR_ARG", -2: "RTRT_ERR_CUDA", -3: "RTRT_ERR_CAPACITY",
                -4: "RTRT_ERR_STATE", -5: "RTRT_ERR_IO", -6: "RTRT_ERR_NCCL"}

# rtrt_buffer
BUF_TRIANGLES, BUF_AABBS, BUF_MORTON_PRESORT, BUF_MORTON, BUF_REORDER_IDX, BUF_BVH_NODES, \
    BUF_TLAS_AABBS, BUF_TLAS_MORTON_PRESORT, BUF_TLAS_MORTON, BUF_TLAS_REORDER_IDX, BUF_TLAS_BVH_NODES, BUF_TRI_PERMUTATION = range(12)
BUF_RENDER_COLOR, BUF_NORMAL, BUF_ALBEDO, BUF_DEPTH, BUF_MOTION_VECTOR, BUF_ACCUMULATION_COLOR, BUF_HISTORY_COLOR, \
    BUF_HISTORY_DEPTH, BUF_NOISE_LEVEL, BUF_NOISE_LEVEL_16, BUF_SKY, BUF_SKY_CDF, BUF_SUN, BUF_SUN_CDF, BUF_ACCUM_RADIANCE = range(32, 47)
BUF_OUTPUT_RGBA8, BUF_COLOR64, BUF_EXPOSURE, BUF_HISTOGRAM = range(47, 51)
BUF_WF_COUNTERS = 100
BUF_WIDE_NODES, BUF_WIDE_TRIS, BUF_WIDE_STAT = 101, 102, 103
MULTI_ROWS, MULTI_SAMPLES = 0, 1
NOALLOC_MAX_RAYS = 65536          # RTRT_NOALLOC_MAX_RAYS (include/rtrt.h)


class Config(C.Structure):
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32),
        ("inputMeshFileName", C.c_char * 256),
        ("inputTextureFileNames", (C.c_char * 256) * 4),
        ("numInputTextureFileNames", C.c_int32),
        ("loadCameraAtInit", C.c_int32),
        ("inputCameraFileName", C.c_char * 256),
        ("cameraSaveFileName", C.c_char * 256),
        ("useDynamicResolution", C.c_int32),
        ("targetFps", C.c_float),
        ("maxWidth", C.c_int32), ("maxHeight", C.c_int32),
        ("minWidth", C.c_int32), ("minHeight", C.c_int32),
    ]


class Timings(C.Structure):
    _fields_ = [("build_bvh_ms", C.c_float), ("traverse_ms", C.c_float), ("pathtrace_ms", C.c_float),
                ("denoise_ms", C.c_float), ("kernel_launches", C.c_uint64), ("rays_traced", C.c_uint64),
                ("node_visits", C.c_uint64), ("leaf_tests", C.c_uint64), ("rays_traced_total", C.c_uint64),
                ("pt_kernel_ms", C.c_float * 6), ("round_node_visits", C.c_uint32 * 4), ("round_leaf_tests", C.c_uint32 * 4),
                ("round_rays", C.c_uint32 * 4), ("wide_build_ms", C.c_float)]


class SkyParams(C.Structure):      # src/settingParams.h:41-45
    _fields_ = [("timeOfDay", C.c_float), ("sunAxisAngle", C.c_float), ("skyScalar", C.c_float),
                ("sunScalar", C.c_float), ("sunAngle", C.c_float)]


class DenoiseParams(C.Structure):  # src/settingParams.h:144-157
    _fields_ = [(n, C.c_float) for n in (
        "local_denoise_sigma_normal", "local_denoise_sigma_depth", "local_denoise_sigma_material",
        "large_denoise_sigma_normal", "large_denoise_sigma_depth", "large_denoise_sigma_material",
        "temporal_denoise_sigma_normal", "temporal_denoise_sigma_depth", "temporal_denoise_sigma_material",
        "noise_threshold_local", "noise_threshold_large")]


class PassSettings(C.Structure):   # src/settingParams.h:87-91
    _fields_ = [(n, C.c_int32) for n in ("enableTemporalDenoising", "enableLocalSpatialFilter",
                                         "enableWideSpatialFilter", "enableTemporalDenoising2")]


class PostprocessParams(C.Structure):   # src/settingParams.h:92-99, 112-116
    _fields_ = [("exposure", C.c_float), ("gain", C.c_float), ("maxWhite", C.c_float), ("gamma", C.c_float),
                ("enablePostProcess", C.c_int32), ("enableAutoExposure", C.c_int32), ("enableSharpening", C.c_int32),
                ("enableToneMapping", C.c_int32)]


class MicrobenchResult(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("l2_read_gbs", "l1_latency_ns", "l2_latency_ns", "dram_latency_ns",
                                         "l1_latency_cycles", "l2_latency_cycles", "dram_latency_cycles", "reserved_")]


# every symbol include/rtrt.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "rtrt_abi_version": (C.c_int, []),
    "rtrt_create": (C.c_int, [C.POINTER(Config), C.c_int, C.POINTER(C.c_void_p)]),
    "rtrt_destroy": (C.c_int, [C.c_void_p]),
    "rtrt_last_error": (C.c_char_p, [C.c_void_p]),
    "rtrt_config_defaults": (C.c_int, [C.POINTER(Config)]),
    "rtrt_load_config_toml": (C.c_int, [C.c_char_p, C.POINTER(Config)]),
    "rtrt_set_mesh": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]),
    "rtrt_set_mesh_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]),
    "rtrt_set_mesh_bin": (C.c_int, [C.c_void_p, C.c_char_p]),
    "rtrt_set_mesh_obj": (C.c_int, [C.c_void_p, C.c_char_p]),
    "rtrt_write_mesh_bin": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int, C.c_int]),
    "rtrt_mesh_file_read": (C.c_int, [C.c_char_p, C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_uint32), C.c_char_p, C.c_size_t]),
    "rtrt_mesh_file_free": (None, [C.c_void_p]),
    "rtrt_update_vertices": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]),
    "rtrt_set_option_i": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int64]),
    "rtrt_build_bvh": (C.c_int, [C.c_void_p]),
    "rtrt_traverse": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_int]),
    "rtrt_traverse_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_int]),
    "rtrt_traverse_host_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "rtrt_get_buffer": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
    "rtrt_read_buffer": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]),
    "rtrt_get_counts": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "rtrt_sync": (C.c_int, [C.c_void_p]),
    "rtrt_join": (C.c_int, [C.c_void_p]),
    "rtrt_read_buffer_async": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]),
    "rtrt_get_timings": (C.c_int, [C.c_void_p, C.POINTER(Timings)]),
    "rtrt_get_stream": (C.c_void_p, [C.c_void_p]),
    "rtrt_microbench": (C.c_int, [C.c_void_p, C.POINTER(MicrobenchResult)]),
    "rtrt_multi_unique_id": (C.c_int, [C.c_void_p]),
    "rtrt_create_multi": (C.c_int, [C.POINTER(Config), C.c_int, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "rtrt_multi_init": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "rtrt_multi_attach_comm": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "rtrt_multi_link_local": (C.c_int, [C.POINTER(C.c_void_p), C.c_int]),
    "rtrt_multi_set_mode": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "rtrt_update_vertices_multi": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]),
    "rtrt_render_frame_multi": (C.c_int, [C.c_void_p]),
    "rtrt_render_progressive_multi": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32]),
    "rtrt_multi_get_band": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "rtrt_multi_plan_rows": (C.c_int, [C.c_int32] * 7 + [C.c_void_p, C.c_int32]),
    "rtrt_set_camera": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]),
    "rtrt_load_camera_file": (C.c_int, [C.c_void_p, C.c_char_p]),
    "rtrt_save_camera_file": (C.c_int, [C.c_void_p, C.c_char_p]),
    "rtrt_set_render_size": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32]),
    "rtrt_dynamic_resolution_step": (C.c_int, [C.POINTER(Config), C.c_int32, C.c_int32, C.c_float, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "rtrt_set_frame_number": (C.c_int, [C.c_void_p, C.c_int32]),
    "rtrt_set_sky_params": (C.c_int, [C.c_void_p, C.POINTER(SkyParams)]),
    "rtrt_get_sun": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "rtrt_set_denoise_params": (C.c_int, [C.c_void_p, C.POINTER(DenoiseParams), C.POINTER(PassSettings)]),
    "rtrt_set_materials": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_uint32]),
    "rtrt_set_texture": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "rtrt_set_row_range": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32]),
    "rtrt_set_data_dir": (C.c_int, [C.c_void_p, C.c_char_p]),
    "rtrt_pathtrace": (C.c_int, [C.c_void_p]),
    "rtrt_denoise": (C.c_int, [C.c_void_p]),
    "rtrt_render_frame": (C.c_int, [C.c_void_p]),
    "rtrt_accumulate": (C.c_int, [C.c_void_p, C.c_int]),
    "rtrt_set_postprocess_params": (C.c_int, [C.c_void_p, C.c_void_p]),
    "rtrt_set_output_size": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32]),
    "rtrt_postprocess": (C.c_int, [C.c_void_p, C.c_float]),
    "rtrt_write_ppm": (C.c_int, [C.c_void_p, C.c_char_p]),
    "rtrt_set_row_strips": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32]),
}

_lib = None


def load():
    """dlopen librtrt.so and bind every declared symbol.  Raises if the library is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: the CUDA extension is not built. Run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback).")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)      # AttributeError if a declared symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
