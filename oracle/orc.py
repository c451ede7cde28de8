"""ctypes wrapper of oracle/_build/liborc.so (the CPU restatement of the reference's hot path).

ORACLE — TEST INFRASTRUCTURE ONLY.
Parity status: stage 1 (LBVH build) and the traversal are PINNED against the reference's own CUDA kernels compiled for
sm_100 (oracle/refcuda -> oracle/_ref/librefcuda.so, tests/test_refcuda_gpu.py: bit-exact at 10 240 / 65 536 / 1 047 552
triangles); sky maps, one PathTrace launch and the denoiser sequence are pinned against the reference kernels within the
tolerances of DESIGN.md §2 (oracle/refcuda/harness_frame.cu, tests/test_refcuda_frame_gpu.py).
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "liborc.so")

TRIANGLE_DTYPE = np.dtype([
    ("v1", "<f4", 3), ("w1", "<f4"), ("v2", "<f4", 3), ("w2", "<f4"), ("v3", "<f4", 3), ("w3", "<f4"),
    ("v4", "<f4", 3), ("w4", "<f4"), ("n1", "<f4", 3), ("u1", "<f4"), ("n2", "<f4", 3), ("u2", "<f4"),
    ("n3", "<f4", 3), ("u3", "<f4"), ("n4", "<f4", 3), ("u4", "<f4")])
AABB_DTYPE = np.dtype([("min", "<f4", 3), ("max", "<f4", 3)])
BVHNODE_DTYPE = np.dtype([
    ("box1maxX", "<f4"), ("box1maxY", "<f4"), ("box1maxZ", "<f4"), ("box2minX", "<f4"),
    ("box1minX", "<f4"), ("box1minY", "<f4"), ("box1minZ", "<f4"), ("box2minY", "<f4"),
    ("box2maxX", "<f4"), ("box2maxY", "<f4"), ("box2maxZ", "<f4"), ("box2minZ", "<f4"),
    ("idxLeft", "<u4"), ("idxRight", "<u4"), ("isLeftLeaf", "<u4"), ("isRightLeaf", "<u4")])
RAY_DTYPE = np.dtype([("orig", "<f4", 3), ("pad16", "<f4"), ("dir", "<f4", 3), ("pad32", "<f4")])
HIT_DTYPE = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("tri", "<i4")])

_lib = None


def build(force=False):
    srcs = [os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith((".cpp", ".h")) or f == "Makefile"]
    if force or not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in srcs):
        subprocess.check_call(["make", "-C", HERE, "-s"] + (["-B"] if force else []))
    return LIB


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        _lib = C.CDLL(LIB)
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def default_threads():
    return max(1, len(os.sched_getaffinity(0)))


class Bvh:
    """Result arrays of orc_build_bvh, same names as the reference's DebugPrintFile dumps (src/bvh.cu)."""
    pass


def build_bvh(pos, nrm, idx, threads=None):
    lib = load()
    pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3)
    nrm = np.ascontiguousarray(nrm, np.float32).reshape(-1, 3)
    idx = np.ascontiguousarray(idx, np.uint32).reshape(-1, 3)
    V, N = pos.shape[0], idx.shape[0]
    B = (N + 1023) // 1024
    Bp = (B + 1023) // 1024 * 1024
    r = Bvh()
    r.num_tris, r.num_batches, r.num_batches_padded = N, B, Bp
    r.triangles = np.zeros(B * 1024, TRIANGLE_DTYPE)
    r.aabbs = np.zeros(B * 1024, AABB_DTYPE)
    r.morton_presort = np.zeros(B * 1024, np.uint32)
    r.morton = np.zeros(B * 1024, np.uint32)
    r.reorder_idx = np.zeros(B * 1024, np.uint32)
    r.nodes = np.zeros(B * 1024, BVHNODE_DTYPE)
    r.tlas_aabbs = np.zeros(Bp, AABB_DTYPE)
    r.tlas_morton_presort = np.zeros(Bp, np.uint32)
    r.tlas_morton = np.zeros(Bp, np.uint32)
    r.tlas_reorder_idx = np.zeros(Bp, np.uint32)
    r.tlas_nodes = np.zeros(Bp, BVHNODE_DTYPE)
    rc = lib.orc_build_bvh(_p(pos), _p(nrm), _p(idx), C.c_uint32(V), C.c_uint32(N),
                           _p(r.triangles), _p(r.aabbs), _p(r.morton_presort), _p(r.morton), _p(r.reorder_idx), _p(r.nodes),
                           _p(r.tlas_aabbs), _p(r.tlas_morton_presort), _p(r.tlas_morton), _p(r.tlas_reorder_idx),
                           _p(r.tlas_nodes), C.c_int(threads or default_threads()))
    if rc:
        raise RuntimeError(f"orc_build_bvh failed: {rc}")
    return r


def traverse(bvh, rays, any_hit=False, stack_size=16, max_loop=1024, threads=None, counters=None):
    lib = load()
    rays = np.ascontiguousarray(rays, RAY_DTYPE)
    hits = np.zeros(rays.shape[0], HIT_DTYPE)
    cnt = np.zeros(2, np.uint64)
    rc = lib.orc_traverse(_p(bvh.triangles), _p(bvh.nodes), _p(bvh.tlas_nodes), C.c_uint32(bvh.num_batches),
                          _p(rays), C.c_uint32(rays.shape[0]), _p(hits), C.c_int(int(any_hit)), C.c_int(stack_size),
                          C.c_int(max_loop), _p(cnt), C.c_int(threads or default_threads()))
    if rc:
        raise RuntimeError(f"orc_traverse failed: {rc}")
    if counters is not None:
        counters[:] = cnt
    return hits


# --------------------------------------------------------------------------------------------------
# stage 2: environment inputs + path trace
# --------------------------------------------------------------------------------------------------
DATA_DIR = os.path.join(os.path.dirname(HERE), "real-time-ray-tracing_b200", "data")


def load_tables():
    bn = np.fromfile(os.path.join(DATA_DIR, "blue_noise_4spp.bin"), np.uint8)
    hosek = np.fromfile(os.path.join(DATA_DIR, "hosek_sky.bin"), np.float32)
    assert bn.size == 65536 + 2 * 131072 and hosek.size == 2460
    return bn, hosek


def sun_dir(time_of_day=0.25, sun_axis_angle=45.0):
    lib = load()
    out = np.zeros(3, np.float32)
    lib.orc_sun_dir(C.c_float(time_of_day), C.c_float(sun_axis_angle), _p(out))
    return out


class Sky:
    pass


def sky_generate(sun_direction, time_of_day=0.25, sun_axis_angle=45.0, sky_scalar=0.01, sun_scalar=0.01, sun_angle=0.6):
    lib = load()
    _, hosek = load_tables()
    s = Sky()
    s.sky = np.zeros((256 * 512, 4), np.float32); s.sky_cdf = np.zeros(256 * 512, np.float32)
    s.sun = np.zeros((32 * 32, 4), np.float32); s.sun_cdf = np.zeros(32 * 32, np.float32)
    params = np.array([time_of_day, sun_axis_angle, sky_scalar, sun_scalar, sun_angle], np.float32)
    cosmax = C.c_float()
    sd = np.ascontiguousarray(sun_direction, np.float32)
    lib.orc_sky_generate(_p(params), _p(sd), _p(hosek), _p(s.sky), _p(s.sky_cdf), _p(s.sun), _p(s.sun_cdf), C.byref(cosmax))
    s.sun_dir = sd; s.cos_theta_max = np.float32(cosmax.value)
    return s


def mip_chain(rgba16, levels=11):
    """2x2 box mip chain with the truncating store of MipmapGen (src/mipgen.cu:121-139); levels past 1x1 alias the last."""
    a = np.ascontiguousarray(rgba16, np.uint16)
    out = [a]
    while (out[-1].shape[0] != 1 or out[-1].shape[1] != 1) and len(out) < levels:
        p = out[-1].astype(np.uint32)
        h, w = p.shape[:2]
        nh, nw = max(1, h // 2), max(1, w // 2)
        ys = np.minimum(np.arange(nh) * 2, h - 1); ys1 = np.minimum(ys + 1, h - 1)
        xs = np.minimum(np.arange(nw) * 2, w - 1); xs1 = np.minimum(xs + 1, w - 1)
        s = p[ys][:, xs] + p[ys][:, xs1] + p[ys1][:, xs] + p[ys1][:, xs1]
        out.append(np.ascontiguousarray((s // 4).astype(np.uint16)))
    while len(out) < levels:
        out.append(out[-1])
    return out


DEFAULT_TEX = [np.array([[[32768, 32768, 32768, 65535]]], np.uint16), np.array([[[32768, 32768, 65535, 65535]]], np.uint16)]

MATERIAL_DTYPE = np.dtype([("albedo", "<f4", 3), ("type", "<u4"), ("F0", "<f4", 3), ("alpha", "<f4")])


def default_materials():
    """src/init.cu:215-251"""
    m = np.zeros(10, MATERIAL_DTYPE)
    m["albedo"] = 0.8; m["type"] = 1; m["F0"] = (0.56, 0.57, 0.58); m["alpha"] = 0.05
    m["type"][0] = 4; m["albedo"][0] = (0.1, 0.2, 0.9)
    m["type"][1] = 2
    m["type"][2] = 4; m["albedo"][2] = (0.9, 0.2, 0.1)
    m["type"][3] = 0; m["albedo"][3] = 0.9
    m["type"][4] = 3; m["albedo"][4] = 0.9
    m["type"][5] = 1
    m["type"][6] = 0
    m["type"][7] = 0; m["albedo"][7] = (0.9, 0.2, 0.1)
    m["type"][8] = 0; m["albedo"][8] = (0.2, 0.9, 0.1)
    m["type"][9] = 0; m["albedo"][9] = (0.1, 0.2, 0.9)
    return m


class _PtArgs(C.Structure):
    _fields_ = [
        ("triangles", C.c_void_p), ("nodes", C.c_void_p), ("tlas", C.c_void_p), ("numBatches", C.c_uint32),
        ("stackSize", C.c_int), ("maxLoop", C.c_int),
        ("materials", C.c_void_p), ("materialsIdx", C.c_void_p), ("numMaterials", C.c_int), ("numMaterialsIdx", C.c_int),
        ("camera", C.c_void_p), ("historyCamera", C.c_void_p),
        ("sky", C.c_void_p), ("skyCdf", C.c_void_p), ("sun", C.c_void_p), ("sunCdf", C.c_void_p),
        ("sunDir", C.c_float * 3), ("sunAngleCosThetaMax", C.c_float),
        ("blueNoise", C.c_void_p),
        ("tex", (C.c_void_p * 11) * 2), ("texW", (C.c_int * 11) * 2), ("texH", (C.c_int * 11) * 2), ("texMaxLod", C.c_int * 2),
        ("frameNum", C.c_int), ("width", C.c_int), ("height", C.c_int),
        ("color", C.c_void_p), ("mask", C.c_void_p), ("normal", C.c_void_p), ("albedo", C.c_void_p), ("depth", C.c_void_p),
        ("motion", C.c_void_p), ("counters", C.c_void_p), ("threads", C.c_int), ("rowBegin", C.c_int), ("rowEnd", C.c_int)]


class Frame:
    pass


def pathtrace(bvh, camera, width, height, frame_num, sky, history_camera=None, materials=None, materials_idx=None,
              textures=None, stack_size=16, max_loop=1024, threads=None, rows=None):
    """One PathTrace launch on the CPU.  Returns float32 G-buffers (values before the half stores)."""
    lib = load()
    bn, _ = load_tables()
    cam = np.ascontiguousarray(camera, np.float32); hcam = np.ascontiguousarray(camera if history_camera is None else history_camera, np.float32)
    assert cam.nbytes == 176 and hcam.nbytes == 176
    mats = np.ascontiguousarray(default_materials() if materials is None else materials, MATERIAL_DTYPE)
    midx = np.ascontiguousarray(np.full(bvh.num_tris, 3, np.int32) if materials_idx is None else materials_idx, np.int32)
    tex = [mip_chain(t) for t in (DEFAULT_TEX if textures is None else textures)]
    f = Frame()
    n = width * height
    f.color = np.zeros((n, 3), np.float32); f.mask = np.zeros(n, np.uint16); f.normal = np.zeros((n, 3), np.float32)
    f.albedo = np.zeros((n, 3), np.float32); f.depth = np.zeros(n, np.float32); f.motion = np.zeros((n, 2), np.float32)
    f.counters = np.zeros(3, np.uint64)
    a = _PtArgs()
    a.triangles, a.nodes, a.tlas, a.numBatches = _p(bvh.triangles).value, _p(bvh.nodes).value, _p(bvh.tlas_nodes).value, bvh.num_batches
    a.stackSize, a.maxLoop = stack_size, max_loop
    a.materials, a.materialsIdx, a.numMaterials, a.numMaterialsIdx = _p(mats).value, _p(midx).value, mats.shape[0], midx.shape[0]
    a.camera, a.historyCamera = _p(cam).value, _p(hcam).value
    a.sky, a.skyCdf, a.sun, a.sunCdf = _p(sky.sky).value, _p(sky.sky_cdf).value, _p(sky.sun).value, _p(sky.sun_cdf).value
    for k in range(3):
        a.sunDir[k] = float(sky.sun_dir[k])
    a.sunAngleCosThetaMax = float(sky.cos_theta_max)
    a.blueNoise = _p(bn).value
    for t in range(2):
        for l in range(11):
            a.tex[t][l] = _p(tex[t][l]).value; a.texW[t][l] = tex[t][l].shape[1]; a.texH[t][l] = tex[t][l].shape[0]
        a.texMaxLod[t] = 10
    a.frameNum, a.width, a.height = frame_num, width, height
    a.color, a.mask, a.normal, a.albedo = _p(f.color).value, _p(f.mask).value, _p(f.normal).value, _p(f.albedo).value
    a.depth, a.motion, a.counters = _p(f.depth).value, _p(f.motion).value, _p(f.counters).value
    a.threads = threads or default_threads()
    a.rowBegin, a.rowEnd = (0, height) if rows is None else rows
    rc = lib.orc_pathtrace(C.byref(a))
    if rc:
        raise RuntimeError(f"orc_pathtrace failed: {rc}")
    return f


# --------------------------------------------------------------------------------------------------
# stage 3: denoise
# --------------------------------------------------------------------------------------------------
COLOR_DTYPE = np.dtype([("rgb", "<f2", 3), ("mask", "<u2")])
DENOISE_DEFAULTS = np.array([100.0, 0.1, 100.0, 100.0, 0.01, 100.0, 100.0, 0.1, 100.0, 0.001, 0.001], np.float32)   # src/settingParams.h:144-157


class DenoiseState:
    """History carried from frame to frame (AccumulationColorBuffer, HistoryColorBuffer, HistoryDepthBuffer).
    `resize` changes the render size of the next frames and keeps the history images at the size they were written
    (historyRenderWidth/Height, src/kernel.cu:83-84,266)."""

    def __init__(self, width, height):
        n = width * height
        self.W, self.H = width, height
        self.accum = np.zeros(n, COLOR_DTYPE); self.history_color = np.zeros(n, COLOR_DTYPE); self.history_depth = np.zeros(n, np.float16)
        self.accum_dim = (width, height); self.history_dim = (width, height)
        self._alloc_noise()

    def _alloc_noise(self):
        self.noise8 = np.zeros(((self.H + 7) // 8) * ((self.W + 7) // 8), np.float16)
        self.noise16 = np.zeros(((self.H + 15) // 16) * ((self.W + 15) // 16), np.float16)

    def resize(self, width, height):
        self.W, self.H = width, height
        self._alloc_noise()


def dynamic_resolution_step(cur_width, delta_time_ms, target_fps=60.0, min_width=640, max_width=3840):
    """RayTracer::UpdateFrame's controller, src/kernel.cu:78-101, in float32 arithmetic: returns (renderWidth, renderHeight)."""
    f32 = np.float32
    dt, fps = f32(delta_time_ms), f32(target_fps)
    w = int(cur_width)
    high, low = f32(1000.0) / (fps - f32(2)), f32(1000.0) / (fps + f32(2))
    if high < dt or low > dt:
        ratio = np.sqrt((f32(1000.0) / fps) / dt, dtype=f32)
        w = int(f32(w) * ratio)                              # int *= float truncates toward zero
    w = w + (-(w % 16) if w % 16 < 8 else 16 - w % 16)
    w = min(max(w, int(min_width)), int(max_width))          # clampi
    return w, (w // 16) * 9


def denoise(state, frame_num, color, normal, depth, motion, albedo, params=None, passes=(1, 1, 1, 1), threads=None):
    """TemporalSpatialDenoising with ping-pong semantics.  color: COLOR_DTYPE[n]; normal/albedo: f2[n,4]; depth: f2[n];
    motion: f2[n,2].  Returns the denoised colour (COLOR_DTYPE) and updates `state` in place."""
    lib = load()
    out = np.ascontiguousarray(color, COLOR_DTYPE).copy()
    normal = np.ascontiguousarray(normal, np.float16); depth = np.ascontiguousarray(depth, np.float16)
    motion = np.ascontiguousarray(motion, np.float16); albedo = np.ascontiguousarray(albedo, np.float16)
    p = np.ascontiguousarray(DENOISE_DEFAULTS if params is None else params, np.float32)
    ps = np.ascontiguousarray(passes, np.int32)
    n = state.W * state.H
    assert out.size == n, "G-buffers must have the state's current render size"
    same = state.accum_dim == (state.W, state.H) and state.history_dim == (state.W, state.H)
    accum_out = state.accum if same else np.zeros(n, COLOR_DTYPE)
    hist_out = state.history_color if same else np.zeros(n, COLOR_DTYPE)
    hdepth_out = state.history_depth if same else np.zeros(n, np.float16)
    if not same:                                             # passes that are switched off leave the old image (and its size) in place
        if not passes[0]: accum_out = state.accum
        if not passes[3]: hist_out = state.history_color; hdepth_out = state.history_depth
    rc = lib.orc_denoise_hist(C.c_int(state.W), C.c_int(state.H), C.c_int(state.accum_dim[0]), C.c_int(state.accum_dim[1]),
                              C.c_int(state.history_dim[0]), C.c_int(state.history_dim[1]), C.c_int(frame_num), _p(p), _p(ps), _p(out),
                              _p(normal), _p(depth), _p(motion), _p(albedo), _p(state.accum), _p(state.history_color), _p(accum_out),
                              _p(hist_out), _p(hdepth_out), _p(state.noise8), _p(state.noise16), C.c_int(threads or default_threads()))
    if passes[0]: state.accum = accum_out; state.accum_dim = (state.W, state.H)
    if passes[3]: state.history_color = hist_out; state.history_depth = hdepth_out; state.history_dim = (state.W, state.H)
    if rc:
        raise RuntimeError(f"orc_denoise failed: {rc}")
    return out
