"""CPU tests (-m "not gpu"): the oracle against known answers and against domain properties.

The reference ships no golden vectors for this path; the oracle is pinned against the reference's own kernels (golden vectors made by
them: tests/test_golden_refkernels_cpu.py; live on the GPU: tests/test_refcuda*_gpu.py).  This file adds
(1) hand-computed known answers of the integer pieces, (2) structural invariants of the hierarchy,
(3) a brute-force closest-hit over all triangles using the same watertight test.
"""
import ctypes as C

import numpy as np
import pytest

from helpers import tree_leaves


def test_morton_known_answers(orc):
    lib = orc.load()
    lib.orc_morton30.restype = C.c_uint32
    lib.orc_morton30.argtypes = [C.c_uint32] * 3
    # src/updateGeometry.cuh:13-27 — x lowest bit, y<<1, z<<2, 10 bits per axis
    assert lib.orc_morton30(0, 0, 0) == 0
    assert lib.orc_morton30(1, 0, 0) == 1
    assert lib.orc_morton30(0, 1, 0) == 2
    assert lib.orc_morton30(0, 0, 1) == 4
    assert lib.orc_morton30(1023, 1023, 1023) == 0x3FFFFFFF
    assert lib.orc_morton30(1023, 0, 0) == 0x09249249
    ref = 0
    for bit in range(10):
        ref |= ((2 >> bit) & 1) << (3 * bit) | ((3 >> bit) & 1) << (3 * bit + 1) | ((1 >> bit) & 1) << (3 * bit + 2)
    assert lib.orc_morton30(2, 3, 1) == ref
    # values above 10 bits are masked (bit 10 of x is dropped; bits 24,25 of the mask pick up x<<16 leftovers)
    rng = np.random.default_rng(0)
    for x, y, z in rng.integers(0, 1024, size=(64, 3)):
        ref = 0
        for bit in range(10):
            ref |= ((int(x) >> bit) & 1) << (3 * bit) | ((int(y) >> bit) & 1) << (3 * bit + 1) | ((int(z) >> bit) & 1) << (3 * bit + 2)
        assert lib.orc_morton30(int(x), int(y), int(z)) == ref


def test_cuda_float_to_uint_semantics(orc):
    lib = orc.load()
    lib.orc_cuda_f2u.restype = C.c_uint32
    lib.orc_cuda_f2u.argtypes = [C.c_float]
    assert lib.orc_cuda_f2u(float("nan")) == 0          # cvt.rzi.u32.f32: NaN -> 0  (flat axis: 0/0)
    assert lib.orc_cuda_f2u(-3.5) == 0
    assert lib.orc_cuda_f2u(1022.999) == 1022
    assert lib.orc_cuda_f2u(5e9) == 0xFFFFFFFF


@pytest.mark.parametrize("n_tris", [1, 2, 3, 5, 1023, 1024, 1025, 2500, 10240])
def test_bvh_structure(orc, pkg, n_tris):
    pos, nrm, idx = pkg.scenes.terrain(64, 80, seed=3)
    idx = idx[:n_tris]
    b = orc.build_bvh(pos, nrm, idx)
    N, B = n_tris, (n_tris + 1023) // 1024
    assert b.num_batches == B
    for k in range(B):
        c = 1024 if k < B - 1 else N - 1024 * (B - 1)
        seg = slice(k * 1024, (k + 1) * 1024)
        keys, ro = b.morton[seg], b.reorder_idx[seg]
        assert np.all(keys[:-1] <= keys[1:]), "keys sorted"
        assert sorted(ro.tolist()) == list(range(1024)), "reorderIdx is a permutation"
        np.testing.assert_array_equal(b.morton_presort[seg][ro], keys)                     # sorted = presort[reorder]
        active = ((c + 3) // 4) * 4
        assert np.all(b.morton_presort[seg][active:] == 0xFFFFFFFF)
        # stability: equal keys keep ascending original index
        same = keys[:-1] == keys[1:]
        assert np.all(ro[:-1][same] < ro[1:][same])
        nodes = b.nodes[seg]
        leaves = tree_leaves(nodes, c)
        if c > 1:
            assert sorted(leaves) == sorted(ro[:c].tolist()), "every sorted leaf appears exactly once"
        # root box = union of the leaf boxes it covers
        if c > 1:
            lb = b.aabbs[seg][np.array(leaves)]
            root_min = np.minimum([nodes["box1minX"][0], nodes["box1minY"][0], nodes["box1minZ"][0]],
                                  [nodes["box2minX"][0], nodes["box2minY"][0], nodes["box2minZ"][0]])
            root_max = np.maximum([nodes["box1maxX"][0], nodes["box1maxY"][0], nodes["box1maxZ"][0]],
                                  [nodes["box2maxX"][0], nodes["box2maxY"][0], nodes["box2maxZ"][0]])
            np.testing.assert_array_equal(root_min, lb["min"].min(axis=0))
            np.testing.assert_array_equal(root_max, lb["max"].max(axis=0))
    tl = tree_leaves(b.tlas_nodes, B)
    assert sorted(tl) == list(range(B)) or (B == 1 and tl == [0])


def test_aabb_padding_and_centroid(orc, pkg):
    # an axis-aligned (flat) triangle: max = min + max(max-min, eps*max)   src/updateGeometry.cuh:176-177
    pos = np.array([[1, 2, 3], [2, 2, 3], [1, 2, 5]], np.float32)
    nrm = np.tile(np.array([[0, 1, 0]], np.float32), (3, 1))
    idx = np.array([[0, 1, 2]], np.uint32)
    b = orc.build_bvh(pos, nrm, idx)
    eps = np.float32(2.0 ** -23)
    np.testing.assert_array_equal(b.aabbs["min"][0], [1, 2, 3])
    np.testing.assert_array_equal(b.aabbs["max"][0], np.float32([2, np.float32(2) + eps * np.float32(2), 5]))
    # N=1: single-leaf node (src/buildBVH.cuh:31-38): both links leaf 0, right box = point at origin
    n0 = b.nodes[0]
    assert (n0["idxLeft"], n0["idxRight"], n0["isLeftLeaf"], n0["isRightLeaf"]) == (0, 0, 1, 1)
    assert n0["box2maxX"] == 0 and n0["box2minZ"] == 0
    # pad items 1..3 use vertex 0 three times (src/init.cu:103-115) and still get AABBs / keys
    np.testing.assert_array_equal(b.aabbs["min"][1], pos[0])
    assert np.all(b.morton_presort[:4] != 0xFFFFFFFF) and np.all(b.morton_presort[4:1024] == 0xFFFFFFFF)


def test_duplicate_keys_give_valid_tree(orc):
    # 1024 identical triangles -> all keys equal: the reference's LCP (no tie-break) breaks; ours must not
    pos = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    nrm = np.tile(np.array([[0, 0, 1]], np.float32), (3, 1))
    idx = np.tile(np.array([[0, 1, 2]], np.uint32), (1024, 1))
    b = orc.build_bvh(pos, nrm, idx)
    assert len(set(b.morton[:1024].tolist())) == 1
    assert sorted(tree_leaves(b.nodes[:1024], 1024)) == list(range(1024))


def _brute_force(orc, bvh, rays):
    lib = orc.load()
    hits = np.zeros(rays.shape[0], orc.HIT_DTYPE)
    lib.orc_brute_force_closest(bvh.triangles.ctypes.data_as(C.c_void_p), C.c_uint32(bvh.num_tris),
                                rays.ctypes.data_as(C.c_void_p), C.c_uint32(rays.shape[0]),
                                hits.ctypes.data_as(C.c_void_p), C.c_int(orc.default_threads()))
    return hits


def test_traversal_matches_brute_force(orc, pkg):
    pos, nrm, idx = pkg.scenes.terrain(64, 80, seed=5)
    b = orc.build_bvh(pos, nrm, idx)
    cam = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, 96, 96)
    rays = pkg.scenes.primary_rays(cam, 96, 96)
    # generous stack / loop so the reference's caps cannot hide a subtree
    got = orc.traverse(b, rays, stack_size=64, max_loop=1 << 20)
    want = _brute_force(orc, b, rays)
    np.testing.assert_array_equal(got["t"].view(np.uint32), want["t"].view(np.uint32))
    hit = want["tri"] >= 0
    assert hit.mean() > 0.3
    # the BVH order may pick a different triangle on exact ties only
    diff = got["tri"] != want["tri"]
    assert diff.sum() <= 0.001 * hit.sum()
    # with the reference caps (16 / 1024) this scene is still fully resolved
    ref_caps = orc.traverse(b, rays)
    np.testing.assert_array_equal(ref_caps["t"].view(np.uint32), got["t"].view(np.uint32))


def test_any_hit_is_consistent(orc, pkg):
    pos, nrm, idx = pkg.scenes.terrain(64, 80, seed=5)
    b = orc.build_bvh(pos, nrm, idx)
    cam = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, 64, 64)
    rays = pkg.scenes.primary_rays(cam, 64, 64)
    closest = orc.traverse(b, rays)
    anyh = orc.traverse(b, rays, any_hit=True)
    np.testing.assert_array_equal(anyh["tri"] >= 0, closest["tri"] >= 0)
    assert np.all(anyh["t"][anyh["tri"] >= 0] >= closest["t"][closest["tri"] >= 0])
