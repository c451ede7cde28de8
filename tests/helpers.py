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
