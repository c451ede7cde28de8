#!/usr/bin/env python
"""Quick per-stage timing on the bench scene (not the bench contract): CUDA events, L2 flushed, median of N."""
import importlib, os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
pkg = importlib.import_module("real-time-ray-tracing_b200")
W, H = (int(os.environ.get("W", 1920)), int(os.environ.get("H", 1080)))
scene = os.environ.get("SCENE", "C2ref")
N = int(os.environ.get("N", 10))
gx, gz = pkg.scenes.TERRAIN_SIZES[scene]
pos, nrm, idx = pkg.scenes.terrain(gx, gz, seed=1, order=os.environ.get("ORDER", "morton"))
cam = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, W, H)
rays = pkg.scenes.primary_rays(cam, W, H)
rt = pkg.RayTracer(W, H).init(); rt.set_mesh(pos, idx, nrm); rt.set_camera(cam, reset_history=True)
for k, v in [a.split("=") for a in sys.argv[1:]]:
    rt.set_option(k, int(v))
stream = torch.cuda.ExternalStream(rt.stream)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
d_rays = torch.from_numpy(rays.view(np.uint8).reshape(-1)).cuda(); d_hits = torch.empty(rays.shape[0] * 16, dtype=torch.uint8, device="cuda")
def timeit(fn, n=N):
    ts = []
    for i in range(n + 2):
        with torch.cuda.stream(stream): flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream); fn(i); b.record(stream); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts[2:]))
out = {"scene": scene, "tris": int(idx.shape[0]), "W": W, "H": H}
out["build_ms"] = timeit(lambda i: rt.build_bvh())
out["traverse_primary_ms"] = timeit(lambda i: rt.traverse_device(d_rays.data_ptr(), rays.shape[0], d_hits.data_ptr()))
def pt(i): rt.set_frame_number(2 + i); rt.pathtrace()
out["pathtrace_ms"] = timeit(pt)
out["rays"] = rt.timings()["rays_traced"]
def dn(i): rt.set_frame_number(2 + i); rt.denoise()
out["denoise_ms"] = timeit(dn)
def pp(i): rt.postprocess(16.0)
out["postprocess_ms"] = timeit(pp)
out["pathtrace_mrays_s"] = out["rays"] / out["pathtrace_ms"] / 1e3
out["primary_mrays_s"] = rays.shape[0] / out["traverse_primary_ms"] / 1e3
try:
    if os.environ.get("NOREF"): raise ImportError("skipped")
    from oracle.refcuda import ref
    if ref.available() and idx.shape[0] < 1023 * 1024 + 1:
        rb = ref.build_bvh(pos, nrm, idx, iterations=6)
        out["ref_build_ms"] = rb.ms
        _, ms = ref.traverse(rb, rays, iterations=6)
        out["ref_traverse_primary_ms"] = ms; out["ref_primary_mrays_s"] = rays.shape[0] / ms / 1e3
        if hasattr(ref.load(), "refcuda_frame_create"):
            from oracle import orc
            orc.build()
            for fm in ([False, True] if ref.available(fmad=True) else [False]):
                tag = "ref_fmad_" if fm else "ref_"
                rbf = ref.build_bvh(pos, nrm, idx, fmad=fm)
                with ref.RefFrame(W, H, fmad=fm) as rf:
                    sky = orc.Sky(); sky.sky = rt.read_buffer(pkg.BUF_SKY); sky.sky_cdf = rt.read_buffer(pkg.BUF_SKY_CDF)
                    sky.sun = rt.read_buffer(pkg.BUF_SUN); sky.sun_cdf = rt.read_buffer(pkg.BUF_SUN_CDF); sky.sun_dir, sky.cos_theta_max = rt.get_sun()
                    rf.set_tables(orc.load_tables()[0]); rf.set_scene(rbf)
                    rf.set_materials(orc.default_materials(), np.full(rbf.num_batches * 1024, 3, np.int32))
                    rf.set_texture(0, orc.mip_chain(orc.DEFAULT_TEX[0])); rf.set_texture(1, orc.mip_chain(orc.DEFAULT_TEX[1])); rf.set_sky(sky)
                    out[tag + "pathtrace_ms"] = rf.pathtrace(cam, 2, iterations=6)
                    out[tag + "pathtrace_mrays_s"] = out["rays"] / out[tag + "pathtrace_ms"] / 1e3
                    dn_ms = []
                    for fr in range(1, 7):
                        rf.pathtrace(cam, fr)
                        dn_ms.append(rf.denoise(fr))
                    out[tag + "denoise_ms"] = float(np.median(dn_ms[2:]))
                    if hasattr(ref.load(), "refcuda_frame_postprocess"):
                        out[tag + "postprocess_ms"] = float(np.median([rf.postprocess(fr, 16.0) for fr in range(1, 7)][2:]))
        if ref.available(fmad=True):
            rb2 = ref.build_bvh(pos, nrm, idx, iterations=6, fmad=True)
            _, ms2 = ref.traverse(rb2, rays, iterations=6, fmad=True)
            out["ref_fmad_build_ms"] = rb2.ms; out["ref_fmad_primary_mrays_s"] = rays.shape[0] / ms2 / 1e3
except ImportError:
    pass
except Exception as e:
    out["ref_error"] = repr(e)
print(json.dumps(out))
