"""GPU parity tests, stage 2 (path trace) and its environment inputs.

Tolerances (stated here, used below):
  * sky map and CDFs (transcendental-heavy, device libm vs glibc): |diff| <= 2e-4 of the map maximum;
    sun-disc map: the limb-darkening argument sqrt(1 - sin^2(gamma)/sin^2(r)) with gamma = acos(dot ~ 1) is
    ill-conditioned (one ulp of the dot product moves the limb texels by ~1e-3), so max |diff| <= 1e-2 and
    mean |diff| <= 2e-4 of the map maximum;
  * G-buffers of one PathTrace launch vs the CPU oracle fed with the SAME sky maps / CDFs / sun direction:
      - material mask, hit/miss: identical on >= 99.9 % of the pixels (a pixel can flip only when a stochastic branch
        `rand < p` sits within float rounding of the device/host libm difference);
      - radiance (demodulated colour), albedo, normal, depth, motion: relative RMSE <= 1e-3 over the pixels whose
        discrete path agrees, after the reference's half-precision store.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

W = H = 256


def rel_rmse(a, b):
    a = a.astype(np.float64); b = b.astype(np.float64)
    return float(np.sqrt(np.mean((a - b) ** 2)) / max(np.sqrt(np.mean(b ** 2)), 1e-12))


def product_sky(rt, pkg, orc):
    s = orc.Sky()
    s.sky = rt.read_buffer(pkg.BUF_SKY); s.sky_cdf = rt.read_buffer(pkg.BUF_SKY_CDF)
    s.sun = rt.read_buffer(pkg.BUF_SUN); s.sun_cdf = rt.read_buffer(pkg.BUF_SUN_CDF)
    s.sun_dir, s.cos_theta_max = rt.get_sun()
    return s


def noise_texture(seed, n=64, normal_map=False):
    rng = np.random.default_rng(seed)
    t = rng.integers(20000, 60000, size=(n, n, 4)).astype(np.uint16)
    if normal_map:   # keep the tangent-space normal mostly +z
        t[..., 0] = rng.integers(30000, 35000, size=(n, n)); t[..., 1] = rng.integers(30000, 35000, size=(n, n)); t[..., 2] = 65535
    return t


@pytest.fixture(scope="module")
def scene(pkg, orc):
    pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C1"], seed=1)
    return pos, nrm, idx, orc.build_bvh(pos, nrm, idx)


def test_sky_maps_match_oracle(pkg, orc):
    with pkg.RayTracer(W, H) as rt:
        for tod, axis in [(0.25, 45.0), (0.1, 30.0), (0.4, 60.0)]:
            rt.set_sky_params(timeOfDay=tod, sunAxisAngle=axis)
            s = product_sky(rt, pkg, orc)
            np.testing.assert_allclose(s.sun_dir, orc.sun_dir(tod, axis), atol=2e-6)
            o = orc.sky_generate(s.sun_dir, tod, axis)
            assert abs(float(s.cos_theta_max) - float(o.cos_theta_max)) < 1e-7
            for got, want, tol in [(s.sky, o.sky, 2e-4), (s.sun, o.sun, 1e-2), (s.sky_cdf, o.sky_cdf, 2e-4), (s.sun_cdf, o.sun_cdf, 2e-4)]:
                assert np.isfinite(got).all()
                assert np.max(np.abs(got - want)) <= tol * np.max(np.abs(want))
                assert np.mean(np.abs(got - want)) <= 2e-4 * np.max(np.abs(want))
            assert np.all(np.diff(s.sky_cdf) >= 0) and np.all(np.diff(s.sun_cdf) >= 0)      # inclusive scan of a non-negative pdf


def compare_frame(rt, pkg, f):
    color = rt.read_buffer(pkg.BUF_RENDER_COLOR)
    mask_ok = color["mask"] == f.mask
    got_depth = rt.read_buffer(pkg.BUF_DEPTH).astype(np.float32)
    want_depth = f.depth.astype(np.float16).astype(np.float32)          # the reference stores depth as half (RayMax -> +inf)
    same_hit = np.isinf(got_depth) == np.isinf(want_depth)
    assert mask_ok.mean() >= 0.999 and same_hit.mean() >= 0.999
    hit = np.isfinite(want_depth) & same_hit
    assert rel_rmse(got_depth[hit], want_depth[hit]) <= 1e-3
    nrm = rt.read_buffer(pkg.BUF_NORMAL).astype(np.float32)[:, :3]
    alb = rt.read_buffer(pkg.BUF_ALBEDO).astype(np.float32)[:, :3]
    mot = rt.read_buffer(pkg.BUF_MOTION_VECTOR).astype(np.float32)
    rgb = color["rgb"].astype(np.float32)
    want_rgb = f.color.astype(np.float16).astype(np.float32)
    # pixels whose discrete path agrees: per-pixel relative difference small; count the rest as flipped branches
    d = np.abs(rgb - want_rgb).max(axis=1) / (np.abs(want_rgb).max(axis=1) + 1e-3)
    agree = (d < 0.05) & mask_ok & same_hit
    assert agree.mean() >= 0.995, f"only {agree.mean():.4f} of the pixels follow the same path"
    assert rel_rmse(rgb[agree], want_rgb[agree]) <= 1e-3
    assert rel_rmse(alb[agree], f.albedo.astype(np.float16).astype(np.float32)[agree]) <= 1e-3
    assert rel_rmse(nrm[agree], f.normal.astype(np.float16).astype(np.float32)[agree]) <= 2e-3
    assert rel_rmse(mot[agree], f.motion.astype(np.float16).astype(np.float32)[agree]) <= 1e-3
    return agree.mean()


def test_pathtrace_default_scene(pkg, orc, scene):
    """All-Lambert reference scene (src/init.cu:262-265), constant textures, frames 1..3, two cameras."""
    pos, nrm, idx, ob = scene
    with pkg.RayTracer(W, H) as rt:
        rt.set_mesh(pos, idx, nrm); rt.build_bvh()
        sky = product_sky(rt, pkg, orc)
        for campos, yaw, pitch in [((0, 30, -60), 0.0, -0.4), ((-2, 12, -2), 0.7, -0.2)]:
            cam = pkg.scenes.make_camera(campos, yaw, pitch, W, H)
            rt.set_camera(cam, reset_history=True)
            for frame in (1, 2, 3):
                rt.set_frame_number(frame)
                rt.pathtrace()
                f = orc.pathtrace(ob, cam, W, H, frame, sky)
                compare_frame(rt, pkg, f)
                rays = rt.timings()["rays_traced"]
                assert rays == int(f.counters[0]) or abs(rays - int(f.counters[0])) <= 0.002 * rays
        assert np.isfinite(rt.read_buffer(pkg.BUF_RENDER_COLOR)["rgb"].astype(np.float32)).all()


def test_pathtrace_mixed_materials_textures_motion(pkg, orc, scene):
    """Lambert / GGX / mirror / glass by triangle id, noise textures (mip chain + bicubic LOD), moving camera (motion vectors)."""
    pos, nrm, idx, ob = scene
    mats = orc.default_materials()
    midx = np.array([3, 4, 5, 1, 7], np.int32)[np.arange(idx.shape[0]) % 5]
    tex = [noise_texture(1), noise_texture(2, normal_map=True)]
    cam0 = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, W, H)
    cam1 = pkg.scenes.make_camera((0.5, 30.2, -59.5), 0.02, -0.41, W, H)
    with pkg.RayTracer(W, H) as rt:
        rt.set_mesh(pos, idx, nrm); rt.build_bvh()
        rt.set_materials(mats.view(pkg.MATERIAL_DTYPE), midx)
        rt.set_texture(0, tex[0]); rt.set_texture(1, tex[1])
        sky = product_sky(rt, pkg, orc)
        rt.set_camera(cam0, reset_history=True); rt.set_frame_number(5); rt.pathtrace()
        compare_frame(rt, pkg, orc.pathtrace(ob, cam0, W, H, 5, sky, materials=mats, materials_idx=midx, textures=tex))
        rt.set_camera(cam1); rt.set_frame_number(6); rt.pathtrace()                      # history camera = cam0
        f = orc.pathtrace(ob, cam1, W, H, 6, sky, history_camera=cam0, materials=mats, materials_idx=midx, textures=tex)
        compare_frame(rt, pkg, f)
        assert np.abs(f.motion - 0.5).max() > 1e-3                                      # the camera did move


def test_any_hit_rays_leave_every_output_unchanged(pkg, orc, scene):
    """Option any_hit_rays (shadow rays and last-segment rays stop at the first accepted triangle): all G-buffers and the ray count
    are BIT-IDENTICAL to the closest-hit search of the reference (any_hit_rays = 0); only the visit counters may shrink.
    Mixed materials (incl. the EMISSIVE and glass types) so that specular chains and the last segment are exercised."""
    pos, nrm, idx, ob = scene
    mats = orc.default_materials()
    midx = np.array([3, 4, 5, 1, 7, 0, 2], np.int32)[np.arange(idx.shape[0]) % 7]
    tex = [noise_texture(1), noise_texture(2, normal_map=True)]
    bufs = (pkg.BUF_RENDER_COLOR, pkg.BUF_NORMAL, pkg.BUF_ALBEDO, pkg.BUF_DEPTH, pkg.BUF_MOTION_VECTOR)
    outs = {}
    for any_hit in (0, 1):
        with pkg.RayTracer(W, H) as rt:
            rt.set_option("any_hit_rays", any_hit); rt.set_option("count_visits", 1)
            rt.set_mesh(pos, idx, nrm); rt.build_bvh()
            rt.set_materials(mats.view(pkg.MATERIAL_DTYPE), midx)
            rt.set_texture(0, tex[0]); rt.set_texture(1, tex[1])
            got = []
            for frame, campos in ((3, (0, 30, -60)), (4, (-2, 12, -2))):
                rt.set_camera(pkg.scenes.make_camera(campos, 0.3 * frame, -0.3, W, H), reset_history=True)
                rt.set_frame_number(frame); rt.pathtrace()
                t = rt.timings()
                got.append(([rt.read_buffer(b).tobytes() for b in bufs], t["rays_traced"], t["node_visits"]))
            outs[any_hit] = got
    for (b0, r0, v0), (b1, r1, v1) in zip(outs[0], outs[1]):
        assert r0 == r1
        assert all(x == y for x, y in zip(b0, b1)), "a frame output changed with any_hit_rays"
        assert v1 <= v0


def test_camera_file_roundtrip(pkg, tmp_path):
    cam = pkg.scenes.make_camera((1, 2, 3), 0.3, -0.1, 640, 480)
    with pkg.RayTracer(640, 480) as rt:
        rt.set_camera(cam)
        rt.SaveCameraToFile(tmp_path / "camera.bin")
        blob = np.fromfile(tmp_path / "camera.bin", np.float32)
        assert blob.nbytes == 176 and np.array_equal(blob, cam)                          # raw sizeof(Camera) dump, src/inputControl.cu:115-131
        rt.LoadCameraFromFile(tmp_path / "camera.bin")


def test_pathtrace_needs_camera_and_bvh(pkg, scene):
    pos, nrm, idx, _ = scene
    with pkg.RayTracer(W, H) as rt:
        rt.set_mesh(pos, idx, nrm)
        with pytest.raises(pkg.RtrtError):
            rt.pathtrace()
        rt.build_bvh()
        with pytest.raises(pkg.RtrtError) as e:
            rt.pathtrace()
        assert e.value.code == -4
