"""ctypes wrapper of oracle/_build/liborc.so (the CPU restatement of the reference's hot path).

ORACLE — TEST INFRASTRUCTURE ONLY; parity unpinned (the reference ships no golden vectors for this path).
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
