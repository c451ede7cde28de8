"""ctypes wrapper of oracle/_ref/librefcuda.so — the REFERENCE'S OWN CUDA kernels (UpdateSceneGeometry, RadixSort, BuildLBVH,
UpdateTLAS, TraverseBvh; PathTrace; the TemporalSpatialDenoising kernels; Sky, SkySun, Scan) compiled for sm_100 by oracle/refcuda/build_ref.py.  TEST / BASELINE INFRASTRUCTURE ONLY.
Needs a GPU; device memory and copies go through torch."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(os.path.dirname(HERE), "_ref", "librefcuda.so")
LIB_FMAD = os.path.join(os.path.dirname(HERE), "_ref", "librefcuda_fmad.so")

_libs = {}


def available(fmad=False):
    return os.path.exists(LIB_FMAD if fmad else LIB)


def load(fmad=False):
    if fmad not in _libs:
        _libs[fmad] = C.CDLL(LIB_FMAD if fmad else LIB)
    return _libs[fmad]


class RefBvh:
    """Device-resident arrays of the reference build (torch uint8 tensors) + host copies on demand."""

    def host(self, name, dtype):
        return self.dev[name].cpu().numpy().view(dtype)


def build_bvh(pos, nrm, idx, iterations=1, fmad=False):
    """BuildBvhLevel1 + BuildBvhLevel2 (src/bvh.cu:7-97) with the reference kernels.  Returns RefBvh; .ms = mean build time."""
    import torch
    lib = load(fmad)
    pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3); nrm = np.ascontiguousarray(nrm, np.float32).reshape(-1, 3)
    idx = np.ascontiguousarray(idx, np.uint32).reshape(-1, 3)
    N = idx.shape[0]
    B = (N + 1023) // 1024
    assert B < 1024, "reference limit: batchCount < 1024 (src/init.cu:126)"
    n4 = (N + 3) // 4 * 4
    idx_p = np.zeros((n4, 3), np.uint32); idx_p[:N] = idx                     # src/init.cu:103-115: indices padded with 0
    tri_count = np.full(B, 1024, np.uint32); tri_count[-1] = N - 1024 * (B - 1)  # src/init.cu:129-130
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()
    z = lambda nbytes: torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
    r = RefBvh()
    r.num_tris, r.num_batches = N, B
    r.dev = {"pos": dev(pos), "nrm": dev(nrm), "idx": dev(idx_p), "triCount": dev(tri_count), "batchCount": dev(np.array([B], np.uint32)),
             "triangles": z(B * 1024 * 128), "aabbs": z(B * 1024 * 24), "morton": z(B * 1024 * 4), "reorderIdx": z(B * 1024 * 4),
             "nodes": z(B * 1024 * 64), "tlasAabbs": z(1024 * 24), "tlasMorton": z(1024 * 4), "tlasReorderIdx": z(1024 * 4), "tlasNodes": z(1024 * 64)}
    p = lambda k: C.c_void_p(r.dev[k].data_ptr())

    def run(iters):
        """one timed pass of the kernel sequence over the same device arrays: mean ms of `iters` builds (the first of several is warm-up)"""
        ms = C.c_float()
        torch.cuda.synchronize()
        rc = lib.refcuda_build_bvh(p("pos"), p("idx"), p("nrm"), C.c_uint(N), p("triangles"), p("aabbs"), p("morton"), p("reorderIdx"), p("nodes"),
                                   p("tlasAabbs"), p("tlasMorton"), p("tlasReorderIdx"), p("tlasNodes"), p("triCount"), p("batchCount"),
                                   C.c_int(iters), C.byref(ms))
        if rc:
            raise RuntimeError(f"refcuda_build_bvh failed: {rc}")
        return ms.value
    r.rebuild = run
    r.ms = run(iterations)
    return r


def traverse(bvh, rays, iterations=1, fmad=False):
    import torch
    lib = load(fmad)
    rays = np.ascontiguousarray(rays)
    n = rays.shape[0]
    d_rays = torch.from_numpy(rays.view(np.uint8).reshape(-1)).cuda()
    d_hits = torch.zeros(n * 16, dtype=torch.uint8, device="cuda")
    ms = C.c_float()
    torch.cuda.synchronize()
    rc = lib.refcuda_traverse(C.c_void_p(bvh.dev["triangles"].data_ptr()), C.c_void_p(bvh.dev["nodes"].data_ptr()),
                              C.c_void_p(bvh.dev["tlasNodes"].data_ptr()), C.c_uint(bvh.num_batches), C.c_void_p(d_rays.data_ptr()), C.c_uint(n),
                              C.c_void_p(d_hits.data_ptr()), C.c_int(iterations), C.byref(ms))
    if rc:
        raise RuntimeError(f"refcuda_traverse failed: {rc}")
    hit_dtype = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("tri", "<i4")])
    return d_hits.cpu().numpy().view(hit_dtype), ms.value


# ---------------------------------------------------------------------------------------------------------------------
# stage 2 / 3 / sky: PathTrace, TemporalSpatialDenoising, Sky + SkySun + Scan (oracle/refcuda/harness_frame.cu)
# ---------------------------------------------------------------------------------------------------------------------
RB_COLOR, RB_ACCUM, RB_HISTORY, RB_NORMAL, RB_DEPTH, RB_HISTORY_DEPTH, RB_MOTION, RB_NOISE, RB_NOISE16, RB_SKY, RB_SUN, RB_ALBEDO, \
    RB_COLOR4, RB_COLOR16, RB_COLOR64, RB_SCALED, RB_OUTPUT = range(17)
COLOR_DTYPE = np.dtype([("rgb", "<f2", 3), ("mask", "<u2")])


class RefFrame:
    """The reference's G-buffer surfaces + scene bundle; methods launch the reference's own kernels."""

    def __init__(self, width, height, fmad=False):
        self.lib = load(fmad)
        self.W, self.H = width, height
        h = C.c_void_p()
        if self.lib.refcuda_frame_create(C.c_int(width), C.c_int(height), C.byref(h)):
            raise RuntimeError("refcuda_frame_create failed")
        self.h = h
        self._keep = []

    def close(self):
        if self.h:
            self.lib.refcuda_frame_destroy(self.h); self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc, what):
        if rc:
            raise RuntimeError(f"{what} failed: {rc}")

    def set_tables(self, blob):
        """blob = the 327 680-byte table file of the product (sobol | scrambling | ranking), see tools/extract_reference_data.py"""
        b = np.ascontiguousarray(blob, np.uint8)
        assert b.size == 65536 + 131072 + 131072
        p = b.ctypes.data
        self._ck(self.lib.refcuda_frame_set_tables(self.h, C.c_void_p(p), C.c_void_p(p + 65536), C.c_void_p(p + 65536 + 131072)), "set_tables")

    def set_scene(self, bvh):
        self._keep.append(bvh)
        self._ck(self.lib.refcuda_frame_set_scene(self.h, C.c_void_p(bvh.dev["triangles"].data_ptr()), C.c_void_p(bvh.dev["nodes"].data_ptr()),
                                                  C.c_void_p(bvh.dev["tlasNodes"].data_ptr()), C.c_int(bvh.num_batches * 1024), C.c_int(bvh.num_batches)),
                 "set_scene")

    def set_materials(self, materials, idx):
        m = np.ascontiguousarray(materials); i = np.ascontiguousarray(idx, np.int32)
        assert m.dtype.itemsize == 32
        self._ck(self.lib.refcuda_frame_set_materials(self.h, C.c_void_p(m.ctypes.data), C.c_int(m.shape[0]), C.c_void_p(i.ctypes.data), C.c_int(i.shape[0])),
                 "set_materials")

    def set_texture(self, slot, mips):
        """mips = list of (h, w, 4) uint16 arrays, level 0 first (orc.mip_chain)"""
        for l, m in enumerate(mips):
            a = np.ascontiguousarray(m, np.uint16)
            self._ck(self.lib.refcuda_frame_set_texture_level(self.h, C.c_int(slot), C.c_int(l), C.c_int(a.shape[1]), C.c_int(a.shape[0]),
                                                              C.c_void_p(a.ctypes.data)), "set_texture_level")
            if a.shape[0] == 1 and a.shape[1] == 1:
                break                                   # the harness aliases the remaining levels to this one

    def generate_sky(self, sun_dir, time_of_day=0.25, sun_axis_angle=45.0, sky_scalar=0.01, sun_scalar=0.01, sun_angle=0.6):
        sd = np.ascontiguousarray(sun_dir, np.float32)
        self._ck(self.lib.refcuda_frame_generate_sky(self.h, C.c_void_p(sd.ctypes.data), C.c_float(time_of_day), C.c_float(sun_axis_angle),
                                                     C.c_float(sky_scalar), C.c_float(sun_scalar), C.c_float(sun_angle)), "generate_sky")

    def set_sky(self, sky):
        """sky: object with .sky (n,4) f32, .sky_cdf, .sun, .sun_cdf, .sun_dir, .cos_theta_max (the oracle's Sky shape)"""
        a = [np.ascontiguousarray(x, np.float32) for x in (sky.sky, sky.sky_cdf, sky.sun, sky.sun_cdf, sky.sun_dir)]
        cosmax = float(sky.cos_theta_max)
        r = np.arccos(np.float32(cosmax))
        sun_area = float(np.tan(r) ** 2 * np.pi)
        self._ck(self.lib.refcuda_frame_set_sky(self.h, *[C.c_void_p(x.ctypes.data) for x in a], C.c_float(cosmax), C.c_float(sun_area)), "set_sky")

    def sky_cdfs(self):
        sky = np.zeros(512 * 256, np.float32); sun = np.zeros(32 * 32, np.float32); c = C.c_float()
        self._ck(self.lib.refcuda_frame_get_sky_cdfs(self.h, C.c_void_p(sky.ctypes.data), C.c_void_p(sun.ctypes.data), C.byref(c)), "get_sky_cdfs")
        return sky, sun, np.float32(c.value)

    def pathtrace(self, camera, frame_num, history_camera=None, iterations=1):
        cam = np.ascontiguousarray(camera, np.float32); hc = cam if history_camera is None else np.ascontiguousarray(history_camera, np.float32)
        assert cam.nbytes == 176 and hc.nbytes == 176
        ms = C.c_float()
        self._ck(self.lib.refcuda_frame_pathtrace(self.h, C.c_void_p(cam.ctypes.data), C.c_void_p(hc.ctypes.data), C.c_int(frame_num),
                                                  C.c_int(iterations), C.byref(ms)), "pathtrace")
        return ms.value

    def denoise(self, frame_num, passes=(1, 1, 1, 1), params=None):
        ps = np.ascontiguousarray(passes, np.int32)
        pp = None if params is None else np.ascontiguousarray(params, np.float32)
        ms = C.c_float()
        self._ck(self.lib.refcuda_frame_denoise(self.h, C.c_int(frame_num), C.c_void_p(ps.ctypes.data),
                                                C.c_void_p(pp.ctypes.data) if pp is not None else None, C.byref(ms)), "denoise")
        return ms.value

    def postprocess(self, frame_num, delta_time_ms=16.0, out_size=None, gain=40.0, max_white=7.0, gamma=2.2):
        """PostProcessing (default passes) + CopyToOutput; returns the elapsed ms.  Read RB_OUTPUT / RB_SCALED / RB_COLOR64 afterwards."""
        ow, oh = out_size or (self.W, self.H)
        self.out_size = (ow, oh)
        pp = np.array([gain, max_white, gamma], np.float32)
        ms = C.c_float()
        self._ck(self.lib.refcuda_frame_postprocess(self.h, C.c_int(frame_num), C.c_float(delta_time_ms), C.c_int(ow), C.c_int(oh),
                                                    C.c_void_p(pp.ctypes.data), C.byref(ms)), "postprocess")
        return ms.value

    def exposure(self):
        e = np.zeros(4, np.float32); hst = np.zeros(64, np.uint32)
        self._ck(self.lib.refcuda_frame_get_exposure(self.h, C.c_void_p(e.ctypes.data), C.c_void_p(hst.ctypes.data)), "get_exposure")
        return e, hst

    _SHAPES = {RB_COLOR4: np.dtype(("<f2", 4)), RB_COLOR16: np.dtype(("<f2", 4)), RB_COLOR64: np.dtype(("<f2", 4)), RB_SCALED: np.dtype(("<f2", 4)),
               RB_OUTPUT: np.dtype(("u1", 4)),
               RB_COLOR: COLOR_DTYPE, RB_ACCUM: COLOR_DTYPE, RB_HISTORY: COLOR_DTYPE, RB_NORMAL: np.dtype(("<f2", 4)), RB_ALBEDO: np.dtype(("<f2", 4)),
               RB_DEPTH: np.dtype("<f2"), RB_HISTORY_DEPTH: np.dtype("<f2"), RB_MOTION: np.dtype(("<f2", 2)), RB_NOISE: np.dtype("<f2"),
               RB_NOISE16: np.dtype("<f2"), RB_SKY: np.dtype(("<f4", 4)), RB_SUN: np.dtype(("<f4", 4))}

    def _count(self, which):
        W, H = self.W, self.H
        c4 = lambda v: (v + 3) // 4
        ow, oh = getattr(self, "out_size", (W, H))
        return {RB_NOISE: ((W + 7) // 8) * ((H + 7) // 8), RB_NOISE16: ((W + 15) // 16) * ((H + 15) // 16), RB_SKY: 512 * 256,
                RB_SUN: 32 * 32, RB_COLOR4: c4(W) * c4(H), RB_COLOR16: c4(c4(W)) * c4(c4(H)), RB_COLOR64: c4(c4(c4(W))) * c4(c4(c4(H))),
                RB_SCALED: ow * oh, RB_OUTPUT: ow * oh}.get(which, W * H)

    def read(self, which):
        dt = self._SHAPES[which]
        out = np.zeros(self._count(which), dt)
        self._ck(self.lib.refcuda_frame_read(self.h, C.c_int(which), C.c_void_p(out.ctypes.data), C.c_size_t(out.nbytes)), "read")
        return out

    def set_render_size(self, w, h):
        """Dynamic resolution: the next frames use the top-left w x h corner of the surfaces (src/kernel.cu:83-101)."""
        self._ck(self.lib.refcuda_frame_set_render_size(self.h, C.c_int(w), C.c_int(h)), "set_render_size")
        self.rW, self.rH = w, h

    def read_region(self, which, w, h):
        out = np.zeros(w * h, self._SHAPES[which])
        self._ck(self.lib.refcuda_frame_read_region(self.h, C.c_int(which), C.c_void_p(out.ctypes.data), C.c_int(w), C.c_int(h)), "read_region")
        return out

    def write_region(self, which, data, w, h):
        a = np.ascontiguousarray(data)
        assert a.nbytes == w * h * np.dtype(self._SHAPES[which]).itemsize
        self._ck(self.lib.refcuda_frame_write_region(self.h, C.c_int(which), C.c_void_p(a.ctypes.data), C.c_int(w), C.c_int(h)), "write_region")

    def write(self, which, data):
        a = np.ascontiguousarray(data)
        self._ck(self.lib.refcuda_frame_write(self.h, C.c_int(which), C.c_void_p(a.ctypes.data), C.c_size_t(a.nbytes)), "write")
