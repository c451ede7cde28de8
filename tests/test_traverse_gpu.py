"""GPU parity tests, stage 2 (traversal): per-ray (triangle id, t, u, v) bit-exact vs the CPU oracle."""
import numpy as np
import pytest

from helpers import assert_hits_equal

pytestmark = pytest.mark.gpu


def _random_rays(pkg, n, seed, lo, hi):
    rng = np.random.default_rng(seed)
    rays = np.zeros(n, pkg.RAY_DTYPE)
    rays["orig"] = rng.uniform(lo, hi, size=(n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    rays["dir"] = d / np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    return rays


def test_config1_primary_rays_256(pkg, orc):
    """BASELINE config 1: ~10k-triangle mesh, LBVH build + primary-ray closest hit at 256x256."""
    pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C1"], seed=1)
    ob = orc.build_bvh(pos, nrm, idx)
    with pkg.RayTracer(256, 256) as rt:
        rt.set_mesh(pos, idx, nrm)
        rt.build_bvh()
        for campos, yaw, pitch in [((0, 30, -60), 0.0, -0.4), ((-2, 2, -2), 0.0, 0.0), ((10, 12, 5), 2.0, -0.9)]:
            cam = pkg.scenes.make_camera(campos, yaw, pitch, 256, 256)
            rays = pkg.scenes.primary_rays(cam, 256, 256)
            got = rt.traverse(rays)
            want = orc.traverse(ob, rays)
            assert_hits_equal(got, want)
        assert (want["tri"] >= 0).any()


def test_random_and_degenerate_rays(pkg, orc):
    pos, nrm, idx = pkg.scenes.terrain(64, 80, seed=2)
    ob = orc.build_bvh(pos, nrm, idx)
    rays = _random_rays(pkg, 50000, 7, (-60, -10, -60), (60, 20, 60))
    # axis-aligned directions (zero components -> SafeDivide eps path, src/linearMath.h:721), rays from inside boxes
    rays["dir"][:300] = np.tile(np.eye(3, dtype=np.float32), (100, 1))
    rays["dir"][300:600] = -np.tile(np.eye(3, dtype=np.float32), (100, 1))
    rays["dir"][600:700] = np.float32([0, -1, 0])
    rays["orig"][600:700, 1] = 40
    with pkg.RayTracer(64, 64) as rt:
        rt.set_mesh(pos, idx, nrm)
        rt.build_bvh()
        assert_hits_equal(rt.traverse(rays), orc.traverse(ob, rays))
        assert_hits_equal(rt.traverse(rays, any_hit=True), orc.traverse(ob, rays, any_hit=True))
        assert rt.traverse(rays[:0]).shape == (0,)          # empty input


def test_stack_and_loop_caps_are_reproduced(pkg, orc):
    """16-entry stack with dropped pushes and the 1024-iteration cap are part of the reference's results."""
    pos, nrm, idx = pkg.scenes.terrain(64, 80, seed=2)
    ob = orc.build_bvh(pos, nrm, idx)
    rays = _random_rays(pkg, 20000, 11, (-60, 0, -60), (60, 6, 60))
    rays["dir"][:, 1] *= 0.05                                # grazing rays: deep stacks, long walks
    rays["dir"] /= np.linalg.norm(rays["dir"], axis=1, keepdims=True)
    with pkg.RayTracer(64, 64) as rt:
        rt.set_mesh(pos, idx, nrm)
        rt.build_bvh()
        for stack, loop in [(16, 1024), (4, 1024), (16, 40), (32, 4096)]:
            rt.set_option("stack_size", stack)
            rt.set_option("max_loop", loop)
            assert_hits_equal(rt.traverse(rays), orc.traverse(ob, rays, stack_size=stack, max_loop=loop))


def test_config2_primary_rays_1080p_1m_tris(pkg, orc):
    """BASELINE config 2 geometry and resolution: 1 047 552 triangles, 1920x1080 primary rays, full compare."""
    gx, gz = pkg.scenes.TERRAIN_SIZES["C2ref"]
    pos, nrm, idx = pkg.scenes.terrain(gx, gz, seed=1)
    ob = orc.build_bvh(pos, nrm, idx)
    cam = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, 1920, 1080)
    rays = pkg.scenes.primary_rays(cam, 1920, 1080)
    with pkg.RayTracer(1920, 1080) as rt:
        rt.set_mesh(pos, idx, nrm)
        rt.build_bvh()
        got = rt.traverse(rays)
    want = orc.traverse(ob, rays)
    assert_hits_equal(got, want)
    assert (want["tri"] >= 0).mean() > 0.3


def test_traverse_before_build_fails(pkg):
    pos, nrm, idx = pkg.scenes.terrain(8, 8)
    with pkg.RayTracer(64, 64) as rt:
        rt.set_mesh(pos, idx, nrm)
        with pytest.raises(pkg.RtrtError) as e:
            rt.traverse(np.zeros(4, pkg.RAY_DTYPE))
        assert e.value.code == -4
