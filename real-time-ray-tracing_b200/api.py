"""Host-side mirror of the reference's `RayTracer` facade (src/kernel.cuh:431-621) over the C ABI.

Method names follow the reference (`init`, `BuildBvhLevel1/2` folded into `build_bvh`, `cleanup`,
`GetBuffer2D` -> `read_buffer`), arguments keep their meaning, errors are raised as `RtrtError`
carrying the rtrt_status code instead of the reference's print-and-exit (src/cudaError.cuh:6-13).
"""
import ctypes as C

import numpy as np

from . import _lib as L

TRIANGLE_DTYPE = np.dtype([  # src/geometry.h:52-97, 128 bytes
    ("v1", "<f4", 3), ("w1", "<f4"), ("v2", "<f4", 3), ("w2", "<f4"), ("v3", "<f4", 3), ("w3", "<f4"),
    ("v4", "<f4", 3), ("w4", "<f4"), ("n1", "<f4", 3), ("u1", "<f4"), ("n2", "<f4", 3), ("u2", "<f4"),
    ("n3", "<f4", 3), ("u3", "<f4"), ("n4", "<f4", 3), ("u4", "<f4")])
AABB_DTYPE = np.dtype([("min", "<f4", 3), ("max", "<f4", 3)])  # src/geometry.h:14-33
BVHNODE_DTYPE = np.dtype([  # src/geometry.h:138-151 + src/bvhNode.cuh:5-13, 64 bytes
    ("box1maxX", "<f4"), ("box1maxY", "<f4"), ("box1maxZ", "<f4"), ("box2minX", "<f4"),
    ("box1minX", "<f4"), ("box1minY", "<f4"), ("box1minZ", "<f4"), ("box2minY", "<f4"),
    ("box2maxX", "<f4"), ("box2maxY", "<f4"), ("box2maxZ", "<f4"), ("box2minZ", "<f4"),
    ("idxLeft", "<u4"), ("idxRight", "<u4"), ("isLeftLeaf", "<u4"), ("isRightLeaf", "<u4")])
RAY_DTYPE = np.dtype([("orig", "<f4", 3), ("pad16", "<f4"), ("dir", "<f4", 3), ("pad32", "<f4")])  # src/geometry.h:40-50
HIT_DTYPE = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("tri", "<i4")])
assert TRIANGLE_DTYPE.itemsize == 128 and AABB_DTYPE.itemsize == 24 and BVHNODE_DTYPE.itemsize == 64
assert RAY_DTYPE.itemsize == 32 and HIT_DTYPE.itemsize == 16

BUFFER_DTYPES = {
    L.BUF_TRIANGLES: TRIANGLE_DTYPE, L.BUF_AABBS: AABB_DTYPE, L.BUF_MORTON_PRESORT: np.dtype("<u4"),
    L.BUF_MORTON: np.dtype("<u4"), L.BUF_REORDER_IDX: np.dtype("<u4"), L.BUF_BVH_NODES: BVHNODE_DTYPE,
    L.BUF_TLAS_AABBS: AABB_DTYPE, L.BUF_TLAS_MORTON_PRESORT: np.dtype("<u4"), L.BUF_TLAS_MORTON: np.dtype("<u4"),
    L.BUF_TLAS_REORDER_IDX: np.dtype("<u4"), L.BUF_TLAS_BVH_NODES: BVHNODE_DTYPE, L.BUF_TRI_PERMUTATION: np.dtype("<u4"),
}

MATERIAL_DTYPE = np.dtype([("albedo", "<f4", 3), ("type", "<u4"), ("F0", "<f4", 3), ("alpha", "<f4")])   # src/kernel.cuh:182-196
COLOR_DTYPE = np.dtype([("rgb", "<f2", 3), ("mask", "<u2")])      # Load2DHalf3Ushort1 layout, src/sampler.cuh:91-128
for _b, _dt in ((L.BUF_RENDER_COLOR, COLOR_DTYPE), (L.BUF_NORMAL, np.dtype(("<f2", 4))), (L.BUF_ALBEDO, np.dtype(("<f2", 4))),
                (L.BUF_DEPTH, np.dtype("<f2")), (L.BUF_MOTION_VECTOR, np.dtype(("<f2", 2))),
                (L.BUF_ACCUMULATION_COLOR, COLOR_DTYPE), (L.BUF_HISTORY_COLOR, COLOR_DTYPE), (L.BUF_HISTORY_DEPTH, np.dtype("<f2")),
                (L.BUF_NOISE_LEVEL, np.dtype("<f2")), (L.BUF_NOISE_LEVEL_16, np.dtype("<f2")),
                (L.BUF_SKY, np.dtype(("<f4", 4))), (L.BUF_SKY_CDF, np.dtype("<f4")), (L.BUF_SUN, np.dtype(("<f4", 4))),
                (L.BUF_SUN_CDF, np.dtype("<f4")), (L.BUF_ACCUM_RADIANCE, np.dtype(("<f4", 4))), (L.BUF_WF_COUNTERS, np.dtype("<u4")),
                (L.BUF_OUTPUT_RGBA8, np.dtype(("u1", 4))), (L.BUF_COLOR64, np.dtype(("<f2", 4))), (L.BUF_EXPOSURE, np.dtype("<f4")),
                (L.BUF_HISTOGRAM, np.dtype("<u4")), (L.BUF_WIDE_NODES, np.dtype(("u1", 128))), (L.BUF_WIDE_TRIS, np.dtype(("<f4", 12))),
                (L.BUF_WIDE_STAT, np.dtype("<u4"))):
    BUFFER_DTYPES[_b] = _dt
MAT_LAMBERTIAN_DIFFUSE, MAT_PERFECT_REFLECTION, MAT_GLASS, MAT_MICROFACET_REFLECTION, MAT_EMISSIVE, MAT_SKY = range(6)

RAY_MAX = np.float32(10e10)


class RtrtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{L.STATUS_NAMES.get(code, code)}: {msg}")
        self.code = code


def load_config(path):
    """LoadConfig (src/configLoader.cpp:5-27) with the path as a parameter."""
    lib = L.load()
    cfg = L.Config()
    rc = lib.rtrt_load_config_toml(str(path).encode(), C.byref(cfg))
    if rc:
        raise RtrtError(rc, f"cannot load {path}")
    return cfg


def config_defaults():
    lib = L.load()
    cfg = L.Config()
    lib.rtrt_config_defaults(C.byref(cfg))
    return cfg


def dynamic_resolution_step(cfg, cur_width, cur_height, delta_time_ms):
    """The dynamic-resolution controller of RayTracer::UpdateFrame (src/kernel.cu:78-101): next (width, height)."""
    w, h = C.c_int32(0), C.c_int32(0)
    rc = L.load().rtrt_dynamic_resolution_step(C.byref(cfg), int(cur_width), int(cur_height), float(delta_time_ms), C.byref(w), C.byref(h))
    if rc:
        raise RtrtError(rc, "rtrt_dynamic_resolution_step: bad argument")
    return w.value, h.value


def read_mesh_file(path):
    """`.obj` (LoadScene, src/fileUtils.cu:16-80) or `.bin` (LoadTrianglesFromFile, src/init.cu:28-50; 128-byte and legacy 64-byte
    triangles) -> (pos [n,3,3] float32, nrm [n,3,3] float32 or None).  Pure host code: works without a GPU."""
    lib = L.load()
    pp, pn, n = C.POINTER(C.c_float)(), C.POINTER(C.c_float)(), C.c_uint32(0)
    err = C.create_string_buffer(256)
    rc = lib.rtrt_mesh_file_read(str(path).encode(), C.byref(pp), C.byref(pn), C.byref(n), err, 256)
    if rc:
        raise RtrtError(rc, f"{path}: {err.value.decode()}")
    try:
        pos = np.ctypeslib.as_array(pp, shape=(n.value, 3, 3)).copy()
        nrm = np.ctypeslib.as_array(pn, shape=(n.value, 3, 3)).copy() if pn else None
    finally:
        lib.rtrt_mesh_file_free(pp)
        if pn:
            lib.rtrt_mesh_file_free(pn)
    return pos, nrm


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class RayTracer:
    """One context per GPU.  `RayTracer(w, h).init()` then `set_mesh` / `build_bvh` / `traverse` ..."""

    def __init__(self, screenWidth=1920, screenHeight=1080, device=0, config=None):
        self._lib = L.load()
        self._h = C.c_void_p()
        self._cfg = config if config is not None else config_defaults()
        if config is None:
            self._cfg.width, self._cfg.height = screenWidth, screenHeight
        self._device = device

    # -- lifetime ---------------------------------------------------------------------------
    def init(self):
        if self._h:
            return self
        rc = self._lib.rtrt_create(C.byref(self._cfg), self._device, C.byref(self._h))
        if rc:
            self._h = C.c_void_p()
            raise RtrtError(rc, "rtrt_create failed (no CUDA device? there is no CPU fallback)")
        return self

    def cleanup(self):
        if self._h:
            self._lib.rtrt_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self.init()

    def __exit__(self, *a):
        self.cleanup()

    def __del__(self):
        try:
            self.cleanup()
        except Exception:
            pass

    def _check(self, rc):
        if rc:
            raise RtrtError(rc, (self._lib.rtrt_last_error(self._h) or b"").decode())

    # -- inputs -----------------------------------------------------------------------------
    def set_mesh(self, pos, idx, nrm=None):
        pos = np.ascontiguousarray(pos, dtype=np.float32).reshape(-1, 3)
        idx = np.ascontiguousarray(idx, dtype=np.uint32).reshape(-1, 3)
        if nrm is not None:
            nrm = np.ascontiguousarray(nrm, dtype=np.float32).reshape(-1, 3)
            assert nrm.shape == pos.shape
        self._check(self._lib.rtrt_set_mesh(self._h, _ptr(pos), _ptr(nrm) if nrm is not None else None, _ptr(idx),
                                            pos.shape[0], idx.shape[0]))
        self._nverts = pos.shape[0]

    def set_mesh_device(self, d_pos, d_nrm, d_idx, num_vertices, num_triangles):
        """Device-resident inputs (ints = device addresses, e.g. torch_tensor.data_ptr()); borrowed."""
        self._check(self._lib.rtrt_set_mesh_device(self._h, C.c_void_p(d_pos), C.c_void_p(d_nrm), C.c_void_p(d_idx),
                                                   num_vertices, num_triangles))

    def update_vertices(self, pos, nrm=None):
        """Per-frame vertex animation (MeshDisplace, src/kernel.cu:139-226): host arrays or raw host addresses."""
        if isinstance(pos, int):
            self._check(self._lib.rtrt_update_vertices(self._h, C.c_void_p(pos), C.c_void_p(nrm) if nrm else None, self._nverts))
            return
        pos = np.ascontiguousarray(pos, dtype=np.float32).reshape(-1, 3)
        if nrm is not None:
            nrm = np.ascontiguousarray(nrm, dtype=np.float32).reshape(-1, 3)
        self._check(self._lib.rtrt_update_vertices(self._h, _ptr(pos), _ptr(nrm) if nrm is not None else None, pos.shape[0]))

    def set_mesh_bin(self, path):
        self._check(self._lib.rtrt_set_mesh_bin(self._h, str(path).encode()))

    def set_mesh_obj(self, path):
        """LoadScene (src/fileUtils.cu:16-80) for Wavefront OBJ: triangles in file order, polygons fanned."""
        self._check(self._lib.rtrt_set_mesh_obj(self._h, str(path).encode()))

    def write_mesh_bin(self, path, presort=True, legacy64=False):
        """tool/meshProcessor.cpp:146-215: the current mesh as `uint32 n; Triangle[n]`, optionally in 60-bit Morton order."""
        self._check(self._lib.rtrt_write_mesh_bin(self._h, str(path).encode(), int(bool(presort)), int(bool(legacy64))))

    def set_option(self, name, value):
        self._check(self._lib.rtrt_set_option_i(self._h, name.encode(), int(value)))

    # -- stages -----------------------------------------------------------------------------
    def build_bvh(self):
        """BuildBvhLevel1 + BuildBvhLevel2 (src/bvh.cu:7-97)."""
        self._check(self._lib.rtrt_build_bvh(self._h))

    def traverse(self, rays, any_hit=False):
        """Host-buffer traversal: rays = structured RAY_DTYPE array -> HIT_DTYPE array."""
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        hits = np.empty(rays.shape[0], dtype=HIT_DTYPE)
        self._check(self._lib.rtrt_traverse_host(self._h, _ptr(rays), rays.shape[0], _ptr(hits), int(any_hit)))
        return hits

    def traverse_noalloc(self, rays, any_hit=False):
        """rtrt_traverse_host_noalloc: the same query with the device staging in the library's static scratch (no cudaMalloc;
        at most L.NOALLOC_MAX_RAYS rays).  Returns (hits, node_visits, leaf_tests)."""
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        hits = np.empty(rays.shape[0], dtype=HIT_DTYPE)
        nv, lt = C.c_uint64(0), C.c_uint64(0)
        self._check(self._lib.rtrt_traverse_host_noalloc(self._h, _ptr(rays), rays.shape[0], _ptr(hits), int(any_hit), C.byref(nv), C.byref(lt)))
        return hits, int(nv.value), int(lt.value)

    def traverse_device(self, d_rays, n, d_hits, any_hit=False):
        self._check(self._lib.rtrt_traverse(self._h, C.c_void_p(d_rays), n, C.c_void_p(d_hits), int(any_hit)))

    # -- frame inputs ------------------------------------------------------------------------
    def set_camera(self, cam, reset_history=False):
        """cam: 44 float32 = the raw Camera struct (src/kernel.cuh:78-101), see scenes.make_camera."""
        cam = np.ascontiguousarray(cam, dtype=np.float32)
        self._check(self._lib.rtrt_set_camera(self._h, _ptr(cam), cam.nbytes, int(reset_history)))

    def LoadCameraFromFile(self, path):
        self._check(self._lib.rtrt_load_camera_file(self._h, str(path).encode()))

    def SaveCameraToFile(self, path):
        self._check(self._lib.rtrt_save_camera_file(self._h, str(path).encode()))

    def set_render_size(self, w, h):
        self._check(self._lib.rtrt_set_render_size(self._h, w, h))

    def set_frame_number(self, n):
        self._check(self._lib.rtrt_set_frame_number(self._h, n))

    def set_sky_params(self, **kw):
        p = L.SkyParams(0.25, 45.0, 0.01, 0.01, 0.6)
        for k, v in kw.items():
            setattr(p, k, v)
        self._check(self._lib.rtrt_set_sky_params(self._h, C.byref(p)))

    def get_sun(self):
        d = (C.c_float * 3)(); c = C.c_float()
        self._check(self._lib.rtrt_get_sun(self._h, d, C.byref(c)))
        return np.array(list(d), np.float32), np.float32(c.value)

    def set_denoise_params(self, params=None, passes=None):
        self._check(self._lib.rtrt_set_denoise_params(self._h, C.byref(params) if params is not None else None,
                                                      C.byref(passes) if passes is not None else None))

    def set_materials(self, materials=None, materials_idx=None):
        m = np.ascontiguousarray(materials, dtype=MATERIAL_DTYPE) if materials is not None else None
        i = np.ascontiguousarray(materials_idx, dtype=np.int32) if materials_idx is not None else None
        self._check(self._lib.rtrt_set_materials(self._h, _ptr(m) if m is not None else None, 0 if m is None else m.shape[0],
                                                 _ptr(i) if i is not None else None, 0 if i is None else i.shape[0]))

    def set_texture(self, slot, rgba16):
        a = np.ascontiguousarray(rgba16, dtype=np.uint16)
        assert a.ndim == 3 and a.shape[2] == 4
        self._check(self._lib.rtrt_set_texture(self._h, slot, a.shape[1], a.shape[0], _ptr(a)))

    def set_row_range(self, begin, end):
        self._check(self._lib.rtrt_set_row_range(self._h, begin, end))

    def set_row_strips(self, strip_height, num_shards, shard):
        """Path trace / accumulate the rows with (row // strip_height) % num_shards == shard (load-balanced multi-GPU split)."""
        self._check(self._lib.rtrt_set_row_strips(self._h, strip_height, num_shards, shard))

    # -- stages 2, 3 ------------------------------------------------------------------------------
    def pathtrace(self):
        """The PathTrace launch of RayTracer::draw (src/kernel.cu:336-357)."""
        self._check(self._lib.rtrt_pathtrace(self._h))

    def denoise(self):
        """TemporalSpatialDenoising (src/denoising.cu:5-188)."""
        self._check(self._lib.rtrt_denoise(self._h))

    def accumulate(self, reset=False):
        """Progressive mode: add colour x albedo of the last path-traced frame to BUF_ACCUM_RADIANCE (rows of the row range)."""
        self._check(self._lib.rtrt_accumulate(self._h, int(reset)))

    def set_postprocess_params(self, params=None):
        self._check(self._lib.rtrt_set_postprocess_params(self._h, C.byref(params) if params is not None else None))

    def set_output_size(self, w, h):
        self._check(self._lib.rtrt_set_output_size(self._h, w, h))

    def postprocess(self, delta_time_ms=16.0):
        """RayTracer::PostProcessing (src/postprocessing.cu:5-160, default passes) + CopyToOutput (src/kernel.cu:26-59) -> BUF_OUTPUT_RGBA8."""
        self._check(self._lib.rtrt_postprocess(self._h, float(delta_time_ms)))

    def write_ppm(self, path):
        self._check(self._lib.rtrt_write_ppm(self._h, str(path).encode()))

    def draw(self):
        """RayTracer::draw for the hot path: BuildBvhLevel1/2 -> PathTrace -> denoise, frameNum++ (src/kernel.cu:259-362)."""
        self._check(self._lib.rtrt_render_frame(self._h))

    # -- observation points -------------------------------------------------------------------
    def counts(self):
        a, b, c = C.c_uint32(), C.c_uint32(), C.c_uint32()
        self._check(self._lib.rtrt_get_counts(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def get_buffer(self, which):
        p, n = C.c_void_p(), C.c_size_t()
        self._check(self._lib.rtrt_get_buffer(self._h, which, C.byref(p), C.byref(n)))
        return p.value, n.value

    def read_buffer(self, which):
        _, nbytes = self.get_buffer(which)
        dt = BUFFER_DTYPES[which]
        out = np.empty(nbytes // dt.itemsize, dtype=dt)
        self._check(self._lib.rtrt_read_buffer(self._h, which, _ptr(out), nbytes))
        return out

    def read_buffer_into(self, which, host_ptr, nbytes):
        """Synchronous copy of a buffer into caller-owned (ideally pinned) host memory at address `host_ptr`."""
        self._check(self._lib.rtrt_read_buffer(self._h, which, C.c_void_p(host_ptr), nbytes))

    def read_buffer_async(self, which, host_ptr, nbytes):
        """Frames in flight: snapshot on the main stream, D2H on the copy stream (overlaps the next frame's kernels).
        The host memory is valid after sync() (or after join() + a wait on the main stream)."""
        self._check(self._lib.rtrt_read_buffer_async(self._h, which, C.c_void_p(host_ptr), nbytes))

    def join(self):
        """Main stream waits for every copy queued on the copy stream (uploads and read-backs)."""
        self._check(self._lib.rtrt_join(self._h))

    def sync(self):
        self._check(self._lib.rtrt_sync(self._h))

    def timings(self):
        t = L.Timings()
        self._check(self._lib.rtrt_get_timings(self._h, C.byref(t)))
        return {k: (list(getattr(t, k)) if hasattr(getattr(t, k), "__len__") else getattr(t, k)) for k, _ in L.Timings._fields_ if k != "reserved_"}

    def microbench(self):
        """rtrt_microbench: measured L2 read bandwidth and L1 / L2 / DRAM dependent-load latencies of this GPU."""
        r = L.MicrobenchResult()
        self._check(self._lib.rtrt_microbench(self._h, C.byref(r)))
        return {k: getattr(r, k) for k, _ in L.MicrobenchResult._fields_ if k != "reserved_"}

    @property
    def stream(self):
        return self._lib.rtrt_get_stream(self._h)
