"""GPU parity tests, stage 1 (LBVH build) — CUDA path through the C ABI vs the CPU oracle, bit-exact on
every observation point of src/bvh.cu (triangles, aabbs, morton pre/post sort, reorderIdx, bvhNodes, TLAS_*)."""
import numpy as np
import pytest

from helpers import assert_bvh_equal, read_all_stage1, tree_leaves

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def rt(pkg):
    r = pkg.RayTracer(256, 256).init()
    r.set_option("keep_presort_morton", 1)
    yield r
    r.cleanup()


@pytest.mark.parametrize("n_tris", [1, 2, 3, 4, 5, 1023, 1024, 1025, 2049, 5000, 10240])
def test_build_matches_oracle_small(rt, pkg, orc, n_tris):
    pos, nrm, idx = pkg.scenes.terrain(64, 80, seed=3)
    idx = idx[:n_tris]
    rt.set_mesh(pos, idx, nrm)
    rt.build_bvh()
    ob = orc.build_bvh(pos, nrm, idx)
    assert rt.counts() == (ob.num_tris, ob.num_batches, ob.num_batches_padded)
    assert_bvh_equal(read_all_stage1(rt, pkg, presort=True), ob, pkg)


def test_build_is_deterministic_and_repeatable(rt, pkg):
    pos, nrm, idx = pkg.scenes.terrain(64, 80, seed=9)
    rt.set_mesh(pos, idx, nrm)
    rt.build_bvh()
    a = read_all_stage1(rt, pkg)
    for _ in range(3):
        rt.build_bvh()           # per-frame rebuild (src/kernel.cu:330-331)
    b = read_all_stage1(rt, pkg)
    for k in a:
        assert a[k].tobytes() == b[k].tobytes()


def test_duplicate_keys(rt, pkg, orc):
    pos = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    nrm = np.tile(np.array([[0, 0, 1]], np.float32), (3, 1))
    idx = np.tile(np.array([[0, 1, 2]], np.uint32), (2048, 1))
    rt.set_mesh(pos, idx, nrm)
    rt.build_bvh()
    ob = orc.build_bvh(pos, nrm, idx)
    bufs = read_all_stage1(rt, pkg, presort=True)
    assert_bvh_equal(bufs, ob, pkg)
    assert sorted(tree_leaves(bufs[pkg.BUF_BVH_NODES][:1024], 1024)) == list(range(1024))


def test_flat_axis_nan_morton(rt, pkg, orc):
    # flat z=0 grid (cf. resources/models/test.bin): (c-min)/(max-min) = 0/0 on z -> cvt.rzi.u32.f32(NaN) = 0
    g = 32
    xs, ys = np.meshgrid(np.linspace(-1, 1, g + 1, dtype=np.float32), np.linspace(-1, 1, g + 1, dtype=np.float32), indexing="ij")
    pos = np.stack([xs, ys, np.zeros_like(xs)], -1).reshape(-1, 3)
    nrm = np.tile(np.array([[0, 0, 1]], np.float32), (pos.shape[0], 1))
    ii, jj = np.meshgrid(np.arange(g), np.arange(g), indexing="ij")
    v00 = ii * (g + 1) + jj
    idx = np.stack([np.stack([v00, v00 + g + 1, v00 + g + 2], -1), np.stack([v00, v00 + g + 2, v00 + 1], -1)], 2).reshape(-1, 3).astype(np.uint32)
    rt.set_mesh(pos, idx, nrm)
    rt.build_bvh()
    ob = orc.build_bvh(pos, nrm, idx)
    assert_bvh_equal(read_all_stage1(rt, pkg, presort=True), ob, pkg)


def test_build_1m_triangles(rt, pkg, orc):
    """BASELINE config 2 geometry: 1 047 552 triangles = 1023 batches (largest the reference accepts)."""
    gx, gz = pkg.scenes.TERRAIN_SIZES["C2ref"]
    pos, nrm, idx = pkg.scenes.terrain(gx, gz, seed=1)
    rt.set_mesh(pos, idx, nrm)
    rt.build_bvh()
    ob = orc.build_bvh(pos, nrm, idx)
    bufs = read_all_stage1(rt, pkg, presort=True)
    assert_bvh_equal(bufs, ob, pkg)
    # size-independent properties
    m = bufs[pkg.BUF_MORTON].reshape(-1, 1024)
    assert np.all(m[:, :-1] <= m[:, 1:])
    ro = np.sort(bufs[pkg.BUF_REORDER_IDX].reshape(-1, 1024), axis=1)
    assert np.all(ro == np.arange(1024)[None, :])
    assert sorted(tree_leaves(bufs[pkg.BUF_TLAS_BVH_NODES], 1023)) == list(range(1023))


def test_device_resident_mesh(pkg, orc):
    import torch
    pos, nrm, idx = pkg.scenes.terrain(64, 80, seed=4)
    dp = torch.from_numpy(pos).cuda(); dn = torch.from_numpy(nrm).cuda(); di = torch.from_numpy(idx.astype(np.int32)).cuda()
    with pkg.RayTracer(64, 64) as r:
        r.set_mesh_device(dp.data_ptr(), dn.data_ptr(), di.data_ptr(), pos.shape[0], idx.shape[0])
        torch.cuda.synchronize()
        r.build_bvh()
        assert_bvh_equal(read_all_stage1(r, pkg), orc.build_bvh(pos, nrm, idx), pkg)
        # animate the vertex buffer in place and rebuild (dynamic geometry, BASELINE config 3 pattern)
        dp[:, 1] += 0.5 * torch.sin(0.3 * dp[:, 0] + 1.0)
        torch.cuda.synchronize()
        r.build_bvh()
        assert_bvh_equal(read_all_stage1(r, pkg), orc.build_bvh(dp.cpu().numpy(), nrm, idx), pkg)


def test_errors(pkg):
    with pkg.RayTracer(64, 64) as r:
        with pytest.raises(pkg.RtrtError) as e:
            r.build_bvh()
        assert e.value.code == -4        # RTRT_ERR_STATE: no mesh
        with pytest.raises(pkg.RtrtError) as e:
            r.set_mesh(np.zeros((3, 3), np.float32), np.array([[0, 1, 7]], np.uint32), np.zeros((3, 3), np.float32))
        assert e.value.code == -1        # index out of range


@pytest.mark.parametrize("quads", [(520, 1024), (1024, 2048)])      # 1 064 960 tris = 1040 batches; 4 194 304 tris = 4096 batches (C3)
def test_build_beyond_1024_batches(pkg, orc, quads):
    """Generalisation rule B7: the reference stops at 1023 batches (src/init.cu:126); the top level then runs the same
    algorithm multi-CTA (device radix sort + global Karras + atomic refit).  Oracle = the same restatement over B items."""
    pos, nrm, idx = pkg.scenes.terrain(*quads, seed=5)
    ob = orc.build_bvh(pos, nrm, idx)
    assert ob.num_batches > 1024
    with pkg.RayTracer(64, 64) as r:
        r.set_option("keep_presort_morton", 1)
        r.set_mesh(pos, idx, nrm)
        r.build_bvh()
        assert r.counts() == (ob.num_tris, ob.num_batches, ob.num_batches_padded)
        bufs = read_all_stage1(r, pkg, presort=True)
        assert_bvh_equal(bufs, ob, pkg)
        assert sorted(tree_leaves(bufs[pkg.BUF_TLAS_BVH_NODES], ob.num_batches)) == list(range(ob.num_batches))
        # and the two-level hierarchy is traversable: primary rays, bit-exact hits
        cam = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, 96, 64)
        rays = pkg.scenes.primary_rays(cam, 96, 64)
        from helpers import assert_hits_equal
        assert_hits_equal(r.traverse(rays), orc.traverse(ob, rays))


@pytest.mark.parametrize("quads,bits", [((64, 80), 60), ((300, 512), 60), ((300, 512), 63)])
def test_global_morton_presort(pkg, orc, quads, bits):
    """option global_sort: the offline CPU pre-sort of tool/meshProcessor.cpp:146-200 (60-bit Morton of the centroid in the mesh
    bounding box, stable) done on the device in front of the batch build.  Shuffled input + device sort must give exactly the
    arrays of host-sorted input + plain build, and the permutation must be the stable argsort of the codes."""
    pos, nrm, idx = pkg.scenes.terrain(*quads, seed=7, order="shuffle")
    c = (pos[idx[:, 0]] + pos[idx[:, 1]] + pos[idx[:, 2]]) / np.float32(3.0)
    code = (pkg.scenes.morton63 if bits == 63 else pkg.scenes.morton60)(c[:, 0], c[:, 1], c[:, 2], pos.min(axis=0), pos.max(axis=0))
    perm = np.argsort(code, kind="stable").astype(np.uint32)
    ob = orc.build_bvh(pos, nrm, idx[perm])
    with pkg.RayTracer(64, 64) as r:
        r.set_option("keep_presort_morton", 1)
        r.set_option("global_sort", 1)
        r.set_option("morton_bits", bits)                 # 63: 21 bits per axis (generalisation rule B7), 60: the reference tool's code
        r.set_mesh(pos, idx, nrm)
        r.build_bvh()
        np.testing.assert_array_equal(r.read_buffer(pkg.BUF_TRI_PERMUTATION), perm)
        assert_bvh_equal(read_all_stage1(r, pkg, presort=True), ob, pkg)
        r.build_bvh()                                     # per-frame re-sort + rebuild is repeatable
        assert_bvh_equal(read_all_stage1(r, pkg, presort=True), ob, pkg)


def test_fused_tlas_timeout_is_an_error(pkg):
    """if the top-level CTA of the fused build gives up waiting for the batch boxes, the build must not report success with a
    stale top level.  tlas_spin_limit = 0 forces the time-out: the next synchronisation point returns RTRT_ERR_CUDA, the BVH is
    invalid, and a normal rebuild recovers."""
    pos, nrm, idx = pkg.scenes.terrain(64, 80, seed=1)
    with pkg.RayTracer(64, 64) as r:
        r.set_mesh(pos, idx, nrm)
        r.set_option("tlas_spin_limit", 0)
        r.build_bvh()
        with pytest.raises(pkg.RtrtError) as e:
            r.sync()
        assert e.value.code == -2
        with pytest.raises(pkg.RtrtError) as e:
            r.traverse(np.zeros(4, pkg.RAY_DTYPE))
        assert e.value.code == -4                      # RTRT_ERR_STATE: BVH not built
        r.set_option("tlas_spin_limit", 1 << 22)
        r.build_bvh(); r.sync()
        assert r.traverse(np.zeros(4, pkg.RAY_DTYPE)).shape == (4,)


@pytest.mark.parametrize("n", [5, 1025, 2050])
def test_keep_all_triangles_option(pkg, n):
    """Reference quirk (src/init.cu:103-115, src/updateGeometry.cuh:93): when N % 4 != 0 the up-to-3 pad items of the last batch get real
    Morton keys and can push real triangles out of the hierarchy.  Default = reproduced; keep_all_triangles = 1 keeps every real one."""
    pos, nrm, idx = pkg.scenes.terrain(64, 80, seed=1)
    idx = np.ascontiguousarray(idx[:n])
    with pkg.RayTracer(64, 64) as r:
        r.set_option("keep_all_triangles", 1)
        r.set_mesh(pos, idx, nrm); r.build_bvh()
        nodes = r.read_buffer(pkg.BUF_BVH_NODES)
        last = n - 1024 * ((n - 1) // 1024)
        leaves = tree_leaves(nodes[1024 * ((n - 1) // 1024):], last)
        assert sorted(leaves) == list(range(last)), "a real triangle of the last batch is missing from the hierarchy"


def test_pathtrace_rejects_deep_parity_stack(pkg):
    pos, nrm, idx = pkg.scenes.terrain(16, 16, seed=1)
    with pkg.RayTracer(64, 64) as r:
        r.set_mesh(pos, idx, nrm); r.build_bvh()
        r.set_camera(pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, 64, 64), reset_history=True)
        r.set_option("stack_size", 32)
        with pytest.raises(pkg.RtrtError) as e:
            r.pathtrace()
        assert e.value.code == -1
        r.set_option("wide_bvh", 1)                     # the wide layout has a spill stack: any stack_size is fine
        r.build_bvh(); r.pathtrace(); r.sync()


def test_contexts_on_two_devices_in_one_process(pkg, orc):
    """kernel attributes (dynamic shared memory opt-in) are per device; contexts on GPUs 0 and 1 of ONE process must both build."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    pos, nrm, idx = pkg.scenes.terrain(64, 80, seed=1)
    ob = orc.build_bvh(pos, nrm, idx)
    cam = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, 64, 64)
    rays = pkg.scenes.primary_rays(cam, 64, 64)
    for dev in (0, 1):
        with pkg.RayTracer(64, 64, device=dev) as r:
            r.set_option("global_sort", 0)
            r.set_mesh(pos, idx, nrm); r.build_bvh()
            assert_bvh_equal(read_all_stage1(r, pkg), ob, pkg)
            from helpers import assert_hits_equal
            assert_hits_equal(r.traverse(rays), orc.traverse(ob, rays))
            r.set_camera(cam, reset_history=True); r.pathtrace(); r.denoise(); r.sync()
