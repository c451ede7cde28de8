#!/usr/bin/env python
"""BASELINE configs 1, 3 and 4 (the parity-test configurations that are not the bench line): product vs the reference's own kernels
(oracle/_ref) vs the CPU oracle, per-stage times with CUDA events, L2 flushed before every timed call.  One JSON object per config."""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
pkg = importlib.import_module("real-time-ray-tracing_b200")
from oracle import orc
orc.build()
try:
    from oracle.refcuda import ref
    HAVE_REF = ref.available() and hasattr(ref.load(), "refcuda_frame_create")
except Exception:
    HAVE_REF = False
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
CORES = orc.default_threads()


def gpu_time(rt, fn, n=10):
    stream = torch.cuda.ExternalStream(rt.stream)
    ts = []
    for i in range(n + 2):
        with torch.cuda.stream(stream):
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream); fn(i); b.record(stream); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts[2:]))


def cpu_time(fn, n=3):
    ts = []
    for _ in range(n):
        t = time.perf_counter(); fn(); ts.append(time.perf_counter() - t)
    return 1e3 * min(ts)


def product_sky(rt):
    s = orc.Sky(); s.sky = rt.read_buffer(pkg.BUF_SKY); s.sky_cdf = rt.read_buffer(pkg.BUF_SKY_CDF)
    s.sun = rt.read_buffer(pkg.BUF_SUN); s.sun_cdf = rt.read_buffer(pkg.BUF_SUN_CDF); s.sun_dir, s.cos_theta_max = rt.get_sun()
    return s


def c1():
    """10 240-tri mesh, LBVH build + 256x256 primary closest hit"""
    W = H = 256
    pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C1"], seed=1)
    cam = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, W, H)
    rays = pkg.scenes.primary_rays(cam, W, H)
    out = {"config": "C1", "tris": int(idx.shape[0]), "rays": int(rays.shape[0]), "cpu_cores": CORES}
    with pkg.RayTracer(W, H) as rt:
        rt.set_mesh(pos, idx, nrm)
        d_rays = torch.from_numpy(rays.view(np.uint8).reshape(-1)).cuda(); d_hits = torch.empty(rays.shape[0] * 16, dtype=torch.uint8, device="cuda")
        out["build_ms"] = gpu_time(rt, lambda i: rt.build_bvh())
        out["traverse_ms"] = gpu_time(rt, lambda i: rt.traverse_device(d_rays.data_ptr(), rays.shape[0], d_hits.data_ptr()))
        got = rt.traverse(rays)
    ob = orc.build_bvh(pos, nrm, idx)
    want = orc.traverse(ob, rays)
    out["bit_exact_vs_oracle"] = bool(all(np.array_equal(got[f].view(np.uint32), want[f].view(np.uint32)) for f in ("tri", "t", "u", "v")))
    out["cpu_build_ms"] = cpu_time(lambda: orc.build_bvh(pos, nrm, idx)); out["cpu_traverse_ms"] = cpu_time(lambda: orc.traverse(ob, rays))
    if HAVE_REF:
        rb = ref.build_bvh(pos, nrm, idx, iterations=11); out["ref_build_ms"] = rb.ms
        rh, ms = ref.traverse(rb, rays, iterations=11); out["ref_traverse_ms"] = ms
        out["bit_exact_vs_reference_kernels"] = bool(all(np.array_equal(got[f].view(np.uint32), rh[f].view(np.uint32)) for f in ("tri", "t", "u", "v")))
    return out


def c3():
    """4 194 304 triangles of dynamic geometry, rebuilt every frame: vertex upload excluded; Morton-ordered input (per-batch sort only)
    and shuffled input (device-side global 60-bit Morton sort + index gather first)"""
    gx, gz = pkg.scenes.TERRAIN_SIZES["C3"]
    out = {"config": "C3", "cpu_cores": CORES}
    for order in ("morton", "shuffle"):
        pos, nrm, idx = pkg.scenes.terrain(gx, gz, seed=1, order=order)
        out["tris"] = int(idx.shape[0])
        with pkg.RayTracer(64, 64) as rt:
            rt.set_mesh(pos, idx, nrm)
            if order == "shuffle":
                rt.set_option("global_sort", 1)
            t = 0.0
            def frame(i):
                rt.build_bvh()
            out[f"build_ms_{order}"] = gpu_time(rt, frame)
        if order == "morton":
            out["cpu_build_ms"] = cpu_time(lambda: orc.build_bvh(pos, nrm, idx), n=2)
    out["algorithmic_GBps_morton"] = 248 * out["tris"] / (out["build_ms_morton"] * 1e-3) / 1e9
    out["algorithmic_GBps_shuffle"] = (248 + 192 + 24) * out["tris"] / (out["build_ms_shuffle"] * 1e-3) / 1e9   # + 63-bit sort passes + index gather
    return out


def c4():
    """temporal + 7x7 + 3 dilated 5x5 + temporal-2 denoise at 3840x2160 on path-traced G-buffers of the C2 scene"""
    W, H = 3840, 2160
    pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C2ref"], seed=1)
    cam = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, W, H)
    out = {"config": "C4", "W": W, "H": H, "cpu_cores": CORES}
    with pkg.RayTracer(W, H) as rt:
        rt.set_mesh(pos, idx, nrm); rt.build_bvh(); rt.set_camera(cam, reset_history=True)
        def pt(i): rt.set_frame_number(2 + i); rt.pathtrace()
        out["pathtrace_ms"] = gpu_time(rt, pt, n=5); out["rays"] = rt.timings()["rays_traced"]
        out["pathtrace_mrays_s"] = out["rays"] / out["pathtrace_ms"] / 1e3
        def dn(i): rt.set_frame_number(20 + i); rt.denoise()
        out["denoise_ms"] = gpu_time(rt, dn)
        out["denoise_GBps_252"] = 252 * W * H / (out["denoise_ms"] * 1e-3) / 1e9
        g = [rt.read_buffer(b) for b in (pkg.BUF_RENDER_COLOR, pkg.BUF_NORMAL, pkg.BUF_DEPTH, pkg.BUF_MOTION_VECTOR, pkg.BUF_ALBEDO)]
        sky = product_sky(rt)
    st = orc.DenoiseState(W, H)
    t = time.perf_counter(); orc.denoise(st, 1, g[0].view(orc.COLOR_DTYPE), g[1], g[2], g[3], g[4])
    t1 = time.perf_counter(); orc.denoise(st, 2, g[0].view(orc.COLOR_DTYPE), g[1], g[2], g[3], g[4]); out["cpu_denoise_ms"] = 1e3 * (time.perf_counter() - t1)
    if HAVE_REF:
        rb = ref.build_bvh(pos, nrm, idx)
        with ref.RefFrame(W, H) as rf:
            rf.set_tables(orc.load_tables()[0]); rf.set_scene(rb); rf.set_materials(orc.default_materials(), np.full(rb.num_batches * 1024, 3, np.int32))
            rf.set_texture(0, orc.mip_chain(orc.DEFAULT_TEX[0])); rf.set_texture(1, orc.mip_chain(orc.DEFAULT_TEX[1])); rf.set_sky(sky)
            out["ref_pathtrace_ms"] = rf.pathtrace(cam, 2, iterations=4)
            dn_ms = []
            for fr in range(1, 7):
                rf.pathtrace(cam, fr); dn_ms.append(rf.denoise(fr))
            out["ref_denoise_ms"] = float(np.median(dn_ms[2:]))
    return out


if __name__ == "__main__":
    which = sys.argv[1:] or ["c1", "c3", "c4"]
    for w in which:
        print(json.dumps({"c1": c1, "c3": c3, "c4": c4}[w]()), flush=True)
