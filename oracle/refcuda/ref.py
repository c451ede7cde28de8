"""ctypes wrapper of oracle/_ref/librefcuda.so — the REFERENCE'S OWN CUDA kernels (UpdateSceneGeometry, RadixSort, BuildLBVH,
UpdateTLAS, TraverseBvh) compiled for sm_100 by oracle/refcuda/build_ref.py.  TEST / BASELINE INFRASTRUCTURE ONLY.
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
    ms = C.c_float()
    torch.cuda.synchronize()
    rc = lib.refcuda_build_bvh(p("pos"), p("idx"), p("nrm"), C.c_uint(N), p("triangles"), p("aabbs"), p("morton"), p("reorderIdx"), p("nodes"),
                               p("tlasAabbs"), p("tlasMorton"), p("tlasReorderIdx"), p("tlasNodes"), p("triCount"), p("batchCount"),
                               C.c_int(iterations), C.byref(ms))
    if rc:
        raise RuntimeError(f"refcuda_build_bvh failed: {rc}")
    r.ms = ms.value
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
