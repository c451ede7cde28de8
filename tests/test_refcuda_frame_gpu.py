"""GPU parity tests of stages 2 and 3 (and the sky inputs) against the REFERENCE'S OWN CUDA kernels: PathTrace
(src/pathtrace.cuh:11-128), the kernels of TemporalSpatialDenoising (src/denoising.cu:48-186, src/temporalDenoising.cuh) and
Sky / SkySun / Scan (src/sky.cuh:280-319, src/scan.cuh), compiled for sm_100 from /root/reference/src by
oracle/refcuda/build_ref.py into oracle/_ref/librefcuda.so and launched by oracle/refcuda/harness_frame.cu.
This pins the product AND the CPU oracle for the floating-point part of the path.  Skipped when the prebuilt library is absent.

Tolerances (stated here, used below; same figures as tests/test_pathtrace_gpu.py and tests/test_denoise_gpu.py):
  * one PathTrace launch on identical inputs (BVH, camera, sky maps + CDFs, blue-noise tables, materials, textures):
    material mask and hit/miss identical on >= 99.9 % of the pixels; radiance, albedo, depth, motion rel-RMSE <= 1e-3 and
    normals <= 2e-3 over the pixels whose discrete path agrees (>= 99.5 %);
  * denoised frame, accumulation and history buffers on identical G-buffers: rel-RMSE <= 2e-2 (the reference filters in place
    while neighbouring blocks read, so it is itself only reproducible to about that figure);
  * sky map / CDFs: <= 2e-4 of the map maximum (sun-disc map: max <= 1e-2, mean <= 2e-4, see tests/test_pathtrace_gpu.py).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

W = H = 256


def rel_rmse(a, b):
    a = a.astype(np.float64); b = b.astype(np.float64)
    return float(np.sqrt(np.mean((a - b) ** 2)) / max(np.sqrt(np.mean(b ** 2)), 1e-12))


@pytest.fixture(scope="module")
def ref():
    from oracle.refcuda import ref as r
    if not r.available():
        pytest.skip("oracle/_ref/librefcuda.so not built (needs /root/reference at build time)")
    if not hasattr(r.load(), "refcuda_frame_create"):
        pytest.skip("oracle/_ref/librefcuda.so predates the frame harness")
    return r


@pytest.fixture(scope="module")
def scene(pkg, orc, ref):
    pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C1"], seed=1)
    return pos, nrm, idx, orc.build_bvh(pos, nrm, idx), ref.build_bvh(pos, nrm, idx)


def product_sky(rt, pkg, orc):
    s = orc.Sky()
    s.sky = rt.read_buffer(pkg.BUF_SKY); s.sky_cdf = rt.read_buffer(pkg.BUF_SKY_CDF)
    s.sun = rt.read_buffer(pkg.BUF_SUN); s.sun_cdf = rt.read_buffer(pkg.BUF_SUN_CDF)
    s.sun_dir, s.cos_theta_max = rt.get_sun()
    return s


def compare_gbuffers(got, want, what):
    """got / want: dicts color (COLOR_DTYPE), normal, albedo (f2 x4), depth (f2), motion (f2 x2)"""
    mask_ok = got["color"]["mask"] == want["color"]["mask"]
    gd, wd = got["depth"].astype(np.float32), want["depth"].astype(np.float32)
    same_hit = np.isinf(gd) == np.isinf(wd)
    assert mask_ok.mean() >= 0.999 and same_hit.mean() >= 0.999, what
    hit = np.isfinite(wd) & same_hit
    assert rel_rmse(gd[hit], wd[hit]) <= 1e-3, what
    rgb, wrgb = got["color"]["rgb"].astype(np.float32), want["color"]["rgb"].astype(np.float32)
    d = np.abs(rgb - wrgb).max(axis=1) / (np.abs(wrgb).max(axis=1) + 1e-3)
    agree = (d < 0.05) & mask_ok & same_hit
    assert agree.mean() >= 0.995, f"{what}: only {agree.mean():.4f} of the pixels follow the same path"
    assert rel_rmse(rgb[agree], wrgb[agree]) <= 1e-3, what
    assert rel_rmse(got["albedo"].astype(np.float32)[agree, :3], want["albedo"].astype(np.float32)[agree, :3]) <= 1e-3, what
    assert rel_rmse(got["normal"].astype(np.float32)[agree, :3], want["normal"].astype(np.float32)[agree, :3]) <= 2e-3, what
    assert rel_rmse(got["motion"].astype(np.float32)[agree], want["motion"].astype(np.float32)[agree]) <= 1e-3, what
    return float(agree.mean()), float((d == 0).mean())


def product_gbuffers(rt, pkg):
    return {"color": rt.read_buffer(pkg.BUF_RENDER_COLOR), "normal": rt.read_buffer(pkg.BUF_NORMAL), "albedo": rt.read_buffer(pkg.BUF_ALBEDO),
            "depth": rt.read_buffer(pkg.BUF_DEPTH), "motion": rt.read_buffer(pkg.BUF_MOTION_VECTOR)}


def ref_gbuffers(rf, ref):
    return {"color": rf.read(ref.RB_COLOR), "normal": rf.read(ref.RB_NORMAL), "albedo": rf.read(ref.RB_ALBEDO),
            "depth": rf.read(ref.RB_DEPTH), "motion": rf.read(ref.RB_MOTION)}


def oracle_gbuffers(f, orc):
    n = f.mask.shape[0]
    color = np.zeros(n, orc.COLOR_DTYPE)
    with np.errstate(over="ignore"):
        color["rgb"] = f.color.astype(np.float16); color["mask"] = f.mask
        nrm = np.zeros((n, 4), np.float16); nrm[:, :3] = f.normal
        alb = np.zeros((n, 4), np.float16); alb[:, :3] = f.albedo
        return {"color": color, "normal": nrm, "albedo": alb, "depth": f.depth.astype(np.float16), "motion": f.motion.astype(np.float16)}


def test_sky_maps_match_reference_kernels(pkg, orc, ref):
    with pkg.RayTracer(W, H) as rt, ref.RefFrame(W, H) as rf:
        for tod, axis in [(0.25, 45.0), (0.1, 30.0), (0.4, 60.0)]:
            rt.set_sky_params(timeOfDay=tod, sunAxisAngle=axis)
            s = product_sky(rt, pkg, orc)
            rf.generate_sky(s.sun_dir, tod, axis)
            sky, sun = rf.read(ref.RB_SKY), rf.read(ref.RB_SUN)
            sky_cdf, sun_cdf, cosmax = rf.sky_cdfs()
            assert abs(float(cosmax) - float(s.cos_theta_max)) < 1e-7
            o = orc.sky_generate(s.sun_dir, tod, axis)
            for name, got, want, tol in [("sky", s.sky, sky, 2e-4), ("sun", s.sun, sun, 1e-2), ("sky_cdf", s.sky_cdf, sky_cdf, 2e-4), ("sun_cdf", s.sun_cdf, sun_cdf, 2e-4),
                                         ("oracle sky", o.sky, sky, 2e-4), ("oracle sun", o.sun, sun, 1e-2), ("oracle sky_cdf", o.sky_cdf, sky_cdf, 2e-4),
                                         ("oracle sun_cdf", o.sun_cdf, sun_cdf, 2e-4)]:
                want = want.reshape(got.shape)
                assert np.isfinite(want).all() and np.max(np.abs(want)) > 0, name
                assert np.max(np.abs(got - want)) <= tol * np.max(np.abs(want)), (name, tod, axis)
                assert np.mean(np.abs(got - want)) <= 2e-4 * np.max(np.abs(want)), (name, tod, axis)


def test_pathtrace_matches_reference_kernel(pkg, orc, ref, scene):
    """Default all-Lambert scene and the mixed-material / textured / moving-camera scene: product and oracle vs PathTrace."""
    pos, nrm, idx, ob, rb = scene
    bn, _ = orc.load_tables()
    mats = orc.default_materials()
    midx_mixed = np.array([3, 4, 5, 1, 7], np.int32)[np.arange(idx.shape[0]) % 5]
    rng = np.random.default_rng(1)
    tex = [rng.integers(20000, 60000, size=(64, 64, 4)).astype(np.uint16), rng.integers(30000, 35000, size=(64, 64, 4)).astype(np.uint16)]
    tex[1][..., 2] = 65535
    cams = [pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, W, H), pkg.scenes.make_camera((0.5, 30.2, -59.5), 0.02, -0.41, W, H)]
    for mixed in (False, True):
        midx = midx_mixed if mixed else np.full(idx.shape[0], 3, np.int32)
        textures = tex if mixed else orc.DEFAULT_TEX
        with pkg.RayTracer(W, H) as rt, ref.RefFrame(W, H) as rf:
            rt.set_mesh(pos, idx, nrm); rt.build_bvh()
            rt.set_materials(mats.view(pkg.MATERIAL_DTYPE), midx)
            if mixed:
                rt.set_texture(0, tex[0]); rt.set_texture(1, tex[1])
            sky = product_sky(rt, pkg, orc)
            rf.set_tables(bn); rf.set_scene(rb); rf.set_materials(mats, np.concatenate([midx, np.full(rb.num_batches * 1024 - idx.shape[0], 3, np.int32)]))
            rf.set_texture(0, orc.mip_chain(textures[0])); rf.set_texture(1, orc.mip_chain(textures[1]))
            rf.set_sky(sky)
            hist = None
            for frame, cam in [(1, cams[0]), (2, cams[0]), (3, cams[1])]:
                rt.set_camera(cam, reset_history=(frame == 1)); rt.set_frame_number(frame)
                rt.pathtrace()
                rf.pathtrace(cam, frame, history_camera=hist)
                want = ref_gbuffers(rf, ref)
                agree, exact = compare_gbuffers(product_gbuffers(rt, pkg), want, f"product vs PathTrace, mixed={mixed}, frame {frame}")
                f = orc.pathtrace(ob, cam, W, H, frame, sky, history_camera=hist, materials=mats, materials_idx=midx, textures=textures)
                compare_gbuffers(oracle_gbuffers(f, orc), want, f"oracle vs PathTrace, mixed={mixed}, frame {frame}")
                assert np.isfinite(want["color"]["rgb"].astype(np.float32)).all()
                hist = cam


@pytest.mark.parametrize("size", [(256, 256), (200, 136)])
def test_denoise_matches_reference_kernels(pkg, orc, ref, scene, size):
    """Five frames with a moving camera: the SAME path-traced G-buffers go through the product's denoiser, the CPU oracle and the
    reference's TemporalFilter / SpatialFilter7x7 / SpatialFilterGlobal5x5 / TemporalFilter2 kernel sequence."""
    Wd, Hd = size
    pos, nrm, idx, ob, rb = scene
    st = orc.DenoiseState(Wd, Hd)
    with pkg.RayTracer(Wd, Hd) as rt, ref.RefFrame(Wd, Hd) as rf:
        rt.set_mesh(pos, idx, nrm); rt.build_bvh()
        for frame in range(1, 6):
            cam = pkg.scenes.make_camera((0.3 * frame, 30, -60 + 0.2 * frame), 0.01 * frame, -0.4, Wd, Hd)
            rt.set_camera(cam, reset_history=(frame == 1)); rt.set_frame_number(frame)
            rt.pathtrace()
            g = product_gbuffers(rt, pkg)
            rf.write(ref.RB_COLOR, g["color"]); rf.write(ref.RB_NORMAL, g["normal"]); rf.write(ref.RB_ALBEDO, g["albedo"])
            rf.write(ref.RB_DEPTH, g["depth"]); rf.write(ref.RB_MOTION, g["motion"])
            rt.denoise()
            rf.denoise(frame)
            o = orc.denoise(st, frame, g["color"].view(orc.COLOR_DTYPE), g["normal"], g["depth"], g["motion"], g["albedo"])
            want = rf.read(ref.RB_COLOR)
            wrgb = want["rgb"].astype(np.float32)
            assert np.isfinite(wrgb).all()
            assert rel_rmse(rt.read_buffer(pkg.BUF_RENDER_COLOR)["rgb"].astype(np.float32), wrgb) <= 2e-2, f"product, frame {frame}"
            assert rel_rmse(o["rgb"].astype(np.float32), wrgb) <= 2e-2, f"oracle, frame {frame}"
            assert np.array_equal(rt.read_buffer(pkg.BUF_RENDER_COLOR)["mask"], want["mask"])
            assert rel_rmse(rt.read_buffer(pkg.BUF_ACCUMULATION_COLOR)["rgb"].astype(np.float32), rf.read(ref.RB_ACCUM)["rgb"].astype(np.float32)) <= 2e-2
            assert rel_rmse(rt.read_buffer(pkg.BUF_HISTORY_COLOR)["rgb"].astype(np.float32), rf.read(ref.RB_HISTORY)["rgb"].astype(np.float32)) <= 2e-2
            np.testing.assert_array_equal(rt.read_buffer(pkg.BUF_HISTORY_DEPTH), rf.read(ref.RB_HISTORY_DEPTH))


def test_denoise_pass_toggles_match_reference_kernels(pkg, orc, ref, scene):
    Wd = Hd = 128
    pos, nrm, idx, ob, rb = scene
    cam = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, Wd, Hd)
    # wide = 2: the 5-iteration a-trous schedule (strides 1, 2, 4, 8, 16) of BASELINE config 4, vs the same reference kernel template
    for passes in [(0, 0, 0, 0), (1, 0, 0, 1), (0, 1, 0, 0), (0, 0, 1, 0), (0, 1, 1, 1), (1, 1, 2, 1), (1, 0, 2, 1)]:
        with pkg.RayTracer(Wd, Hd) as rt, ref.RefFrame(Wd, Hd) as rf:
            rt.set_mesh(pos, idx, nrm); rt.build_bvh()
            rt.set_denoise_params(None, pkg.PassSettings(*passes))
            for frame in (1, 2, 3):
                rt.set_camera(cam, reset_history=(frame == 1)); rt.set_frame_number(frame)
                rt.pathtrace()
                g = product_gbuffers(rt, pkg)
                rf.write(ref.RB_COLOR, g["color"]); rf.write(ref.RB_NORMAL, g["normal"]); rf.write(ref.RB_ALBEDO, g["albedo"])
                rf.write(ref.RB_DEPTH, g["depth"]); rf.write(ref.RB_MOTION, g["motion"])
                rt.denoise(); rf.denoise(frame, passes=passes)
                assert rel_rmse(rt.read_buffer(pkg.BUF_RENDER_COLOR)["rgb"].astype(np.float32), rf.read(ref.RB_COLOR)["rgb"].astype(np.float32)) <= 2e-2, (passes, frame)


def test_dynamic_resolution_history_rescale_matches_reference_kernels(pkg, orc, ref, scene):
    """the render size changes between frames (src/kernel.cu:78-101) while the temporal history keeps the size it was
    written at; TemporalFilter / TemporalFilter2 sample it with historyDim (src/temporalDenoising.cuh:812,837,1028,1053).
    The reference keeps its surfaces at the maximum size and filters the top-left corner, so near the right / bottom edge of a
    shrunk frame its spatial filters and bicubic history taps read stale texels of older, larger frames (surface clamp addressing
    clamps to the SURFACE, not to bufferDim); the product and the oracle clamp to the image.  Interior pixels are compared."""
    maxW, maxH = 640, 360
    sizes = [(640, 360), (640, 360), (480, 270), (480, 270), (560, 315), (640, 360), (640, 360)]
    pos, nrm, idx, ob, rb = scene
    st = orc.DenoiseState(maxW, maxH)
    margin = 64
    with pkg.RayTracer(maxW, maxH) as rt, ref.RefFrame(maxW, maxH) as rf:
        rt.set_mesh(pos, idx, nrm); rt.build_bvh()
        shrunk = False
        for frame, (w, h) in enumerate(sizes, start=1):
            rt.set_render_size(w, h); rf.set_render_size(w, h); st.resize(w, h)
            cam = pkg.scenes.make_camera((0.3 * frame, 30, -60 + 0.2 * frame), 0.01 * frame, -0.4, w, h)
            rt.set_camera(cam, reset_history=(frame == 1)); rt.set_frame_number(frame)
            rt.pathtrace()
            g = product_gbuffers(rt, pkg)
            assert g["color"].size == w * h
            for which, key in [(ref.RB_COLOR, "color"), (ref.RB_NORMAL, "normal"), (ref.RB_ALBEDO, "albedo"), (ref.RB_DEPTH, "depth"), (ref.RB_MOTION, "motion")]:
                rf.write_region(which, g[key], w, h)
            rt.denoise(); rf.denoise(frame)
            o = orc.denoise(st, frame, g["color"].view(orc.COLOR_DTYPE), g["normal"], g["depth"], g["motion"], g["albedo"])
            shrunk = shrunk or (w, h) != (maxW, maxH)
            m = margin if shrunk else 0
            crop = lambda a: a.reshape(h, w, -1)[: h - m, : w - m].astype(np.float32)
            for name, pb, rb_ in [("colour", pkg.BUF_RENDER_COLOR, ref.RB_COLOR), ("accum", pkg.BUF_ACCUMULATION_COLOR, ref.RB_ACCUM),
                                  ("history", pkg.BUF_HISTORY_COLOR, ref.RB_HISTORY)]:
                got = rt.read_buffer(pb); want = rf.read_region(rb_, w, h)
                assert got.size == w * h
                assert np.isfinite(want["rgb"].astype(np.float32)).all()
                assert rel_rmse(crop(got["rgb"]), crop(want["rgb"])) <= 2e-2, f"product {name}, frame {frame} at {w}x{h}"
            assert rel_rmse(crop(o["rgb"]), crop(rf.read_region(ref.RB_COLOR, w, h)["rgb"])) <= 2e-2, f"oracle, frame {frame} at {w}x{h}"
            # product and oracle clamp alike: they must agree over the whole image, edges included
            assert rel_rmse(rt.read_buffer(pkg.BUF_RENDER_COLOR)["rgb"].astype(np.float32), o["rgb"].astype(np.float32)) <= 2e-2, f"product vs oracle, frame {frame}"
            np.testing.assert_array_equal(rt.read_buffer(pkg.BUF_HISTORY_DEPTH), rf.read_region(ref.RB_HISTORY_DEPTH, w, h))


def test_render_size_change_keeps_history(pkg, scene):
    """A size change inside the allocated capacity must not reset the temporal history (the old behaviour re-allocated and zeroed)."""
    pos, nrm, idx, ob, rb = scene
    out = {}
    for keep in (True, False):
        with pkg.RayTracer(320, 180) as rt:
            rt.set_mesh(pos, idx, nrm); rt.build_bvh()
            for frame, (w, h) in enumerate([(320, 180), (320, 180), (256, 144)], start=1):
                rt.set_render_size(w, h)
                rt.set_camera(pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, w, h), reset_history=(frame == 1)); rt.set_frame_number(frame)
                rt.pathtrace()
                if frame == 3 and not keep:
                    rt.set_frame_number(1)           # frameNum 1 = "no history" (src/denoising.cu:50)
                rt.denoise()
            out[keep] = rt.read_buffer(pkg.BUF_RENDER_COLOR)["rgb"].astype(np.float32)
    assert out[True].shape == (256 * 144, 3)
    assert rel_rmse(out[True], out[False]) > 1e-2        # the history did contribute to the resized frame
