"""Shared checks for the parity tests."""
import numpy as np

NODE_BOX_FIELDS = ["box1maxX", "box1maxY", "box1maxZ", "box2minX", "box1minX", "box1minY", "box1minZ", "box2minY",
                   "box2maxX", "box2maxY", "box2maxZ", "box2minZ"]
NODE_LINK_FIELDS = ["idxLeft", "idxRight", "isLeftLeaf", "isRightLeaf"]


def assert_nodes_equal(got, want, count, what):
    """Bit-exact links, value-exact boxes for the first `count` nodes."""
    for f in NODE_LINK_FIELDS:
        np.testing.assert_array_equal(got[f][:count], want[f][:count], err_msg=f"{what}.{f}")
    for f in NODE_BOX_FIELDS:
        np.testing.assert_array_equal(got[f][:count], want[f][:count], err_msg=f"{what}.{f}")


def assert_bvh_equal(rt_bufs, ob, pkg):
    """Compare every stage-1 observation point (src/bvh.cu dumps) of the CUDA build with the oracle's."""
    N, B = ob.num_tris, ob.num_batches
    tri, otri = rt_bufs[pkg.BUF_TRIANGLES], ob.triangles
    last = N - 1024 * (B - 1)
    active = np.ones(B * 1024, bool)
    active[1024 * (B - 1) + ((last + 3) // 4) * 4:] = False          # items the reference's threads never touch
    for f in ["v1", "v2", "v3", "n1", "n2", "n3"]:
        np.testing.assert_array_equal(tri[f][active], otri[f][active], err_msg=f"triangles.{f}")
    for f in ["min", "max"]:
        np.testing.assert_array_equal(rt_bufs[pkg.BUF_AABBS][f][active], ob.aabbs[f][active], err_msg=f"aabbs.{f}")
    np.testing.assert_array_equal(rt_bufs[pkg.BUF_MORTON], ob.morton, err_msg="morton (sorted)")
    np.testing.assert_array_equal(rt_bufs[pkg.BUF_REORDER_IDX], ob.reorder_idx, err_msg="reorderIdx")
    if pkg.BUF_MORTON_PRESORT in rt_bufs:
        np.testing.assert_array_equal(rt_bufs[pkg.BUF_MORTON_PRESORT], ob.morton_presort, err_msg="morton (pre-sort)")
    nodes, onodes = rt_bufs[pkg.BUF_BVH_NODES], ob.nodes
    valid = np.zeros(B * 1024, bool)
    for b in range(B):
        c = 1024 if b < B - 1 else last
        valid[b * 1024: b * 1024 + max(c - 1, 1)] = True
    for f in NODE_LINK_FIELDS + NODE_BOX_FIELDS:
        np.testing.assert_array_equal(nodes[f][valid], onodes[f][valid], err_msg=f"bvhNodes.{f}")
    np.testing.assert_array_equal(rt_bufs[pkg.BUF_TLAS_MORTON], ob.tlas_morton, err_msg="TLAS morton")
    np.testing.assert_array_equal(rt_bufs[pkg.BUF_TLAS_REORDER_IDX], ob.tlas_reorder_idx, err_msg="TLAS reorderIdx")
    for f in ["min", "max"]:
        np.testing.assert_array_equal(rt_bufs[pkg.BUF_TLAS_AABBS][f][:B], ob.tlas_aabbs[f][:B], err_msg=f"TLAS aabbs.{f}")
    assert_nodes_equal(rt_bufs[pkg.BUF_TLAS_BVH_NODES], ob.tlas_nodes, max(B - 1, 1), "TLAS nodes")


def read_all_stage1(rt, pkg, presort=False):
    ids = [pkg.BUF_TRIANGLES, pkg.BUF_AABBS, pkg.BUF_MORTON, pkg.BUF_REORDER_IDX, pkg.BUF_BVH_NODES,
           pkg.BUF_TLAS_AABBS, pkg.BUF_TLAS_MORTON, pkg.BUF_TLAS_REORDER_IDX, pkg.BUF_TLAS_BVH_NODES]
    if presort:
        ids += [pkg.BUF_MORTON_PRESORT, pkg.BUF_TLAS_MORTON_PRESORT]
    return {i: rt.read_buffer(i) for i in ids}


def assert_hits_equal(got, want):
    """Bit-exact per-ray (tri, t, u, v)."""
    np.testing.assert_array_equal(got["tri"], want["tri"], err_msg="hit triangle id")
    np.testing.assert_array_equal(got["t"].view(np.uint32), want["t"].view(np.uint32), err_msg="hit t (bits)")
    np.testing.assert_array_equal(got["u"].view(np.uint32), want["u"].view(np.uint32), err_msg="hit u (bits)")
    np.testing.assert_array_equal(got["v"].view(np.uint32), want["v"].view(np.uint32), err_msg="hit v (bits)")


def tree_leaves(nodes, n):
    """Walk a segment's hierarchy from node 0; returns the list of leaf payloads (iterative DFS)."""
    if n == 1:
        return [int(nodes["idxLeft"][0])]
    out, stack = [], [0]
    visited = 0
    while stack:
        i = stack.pop()
        visited += 1
        assert visited <= n, "cycle in hierarchy"
        for idx, leaf in ((nodes["idxLeft"][i], nodes["isLeftLeaf"][i]), (nodes["idxRight"][i], nodes["isRightLeaf"][i])):
            if leaf:
                out.append(int(idx))
            else:
                stack.append(int(idx))
    assert visited == n - 1
    return out


# ---------------------------------------------------------------------------------------------------------------------
# Golden vectors made by the reference's own kernels (tools/make_golden_refkernels.py -> tests/golden/refkernels_c1.npz):
# the inputs are regenerated from these seeds on both sides, only the reference's outputs are stored.
# ---------------------------------------------------------------------------------------------------------------------
GOLDEN_W = GOLDEN_H = 48


def golden_refkernels_inputs(pkg):
    """BASELINE config 1 in miniature: the 10 240-triangle terrain, 64x48 primary rays + 1 000 incoherent rays, three 48x48 frames
    (frames 1 and 2 from one camera, frame 3 from a moved one, each with the previous frame's camera as history camera)."""
    pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C1"], seed=1)
    rays = pkg.scenes.primary_rays(pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, 64, 48), 64, 48)
    rng = np.random.default_rng(7)
    extra = np.zeros(1000, rays.dtype)
    extra["orig"] = np.stack([rng.uniform(-50, 50, 1000), rng.uniform(5, 30, 1000), rng.uniform(-50, 50, 1000)], -1).astype(np.float32)
    d = rng.normal(size=(1000, 3)).astype(np.float32); d /= np.linalg.norm(d, axis=1, keepdims=True)
    extra["dir"] = d
    cams = [pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, GOLDEN_W, GOLDEN_H), pkg.scenes.make_camera((0.5, 30.2, -59.5), 0.02, -0.41, GOLDEN_W, GOLDEN_H)]
    return {"pos": pos, "nrm": nrm, "idx": idx, "rays": np.concatenate([rays, extra]), "frames": [(1, cams[0]), (2, cams[0]), (3, cams[1])]}


def golden_stage1_digests(b, node_dtype_fields=None):
    """SHA-1 per observation point of a stage-1 result `b` (oracle Bvh or an object of the same shape): sorted Morton codes, reorder
    indices, triangle vertices / normals, BLAS nodes (the 1 023 real nodes of every batch), top-level keys, order, boxes (bits) and nodes."""
    import hashlib
    B = b.num_batches
    valid = np.zeros(B * 1024, bool)
    for k in range(B):
        valid[k * 1024: k * 1024 + 1023] = True
    sha = lambda a: hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()
    out = {"morton": sha(b.morton), "reorder_idx": sha(b.reorder_idx), "tlas_morton": sha(b.tlas_morton[:B]), "tlas_reorder_idx": sha(b.tlas_reorder_idx[:B])}
    for f in ["v1", "v2", "v3", "n1", "n2", "n3"]:
        out["triangles." + f] = sha(b.triangles[f])
    for f in NODE_LINK_FIELDS + NODE_BOX_FIELDS:
        out["nodes." + f] = sha(b.nodes[f][valid])
        out["tlas_nodes." + f] = sha(b.tlas_nodes[f][:B - 1])
    for f in ["min", "max"]:
        out["tlas_aabbs." + f] = sha(b.tlas_aabbs[f][:B])
    return out
