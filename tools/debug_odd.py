#!/usr/bin/env python
"""Where do non-finite denoised pixels come from at odd frame sizes?  (debug helper)"""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
pkg = importlib.import_module("real-time-ray-tracing_b200")
from oracle import orc
orc.build()
for (W, H) in [(201, 136), (200, 137), (201, 137), (208, 137), (193, 129)]:
    pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C1"], seed=1)
    st = orc.DenoiseState(W, H)
    with pkg.RayTracer(W, H) as rt:
        rt.set_mesh(pos, idx, nrm); rt.build_bvh()
        for frame in range(1, 4):
            cam = pkg.scenes.make_camera((0.3 * frame, 30, -60 + 0.2 * frame), 0.01 * frame, -0.4, W, H)
            rt.set_camera(cam, reset_history=(frame == 1)); rt.set_frame_number(frame)
            rt.pathtrace()
            bufs = {n: rt.read_buffer(getattr(pkg, n)) for n in ("BUF_RENDER_COLOR", "BUF_NORMAL", "BUF_DEPTH", "BUF_MOTION_VECTOR", "BUF_ALBEDO")}
            fin_in = {n: bool(np.isfinite((b["rgb"] if b.dtype.names else b).astype(np.float32)).all()) for n, b in bufs.items() if n != "BUF_DEPTH"}
            rt.denoise()
            got = rt.read_buffer(pkg.BUF_RENDER_COLOR)["rgb"].astype(np.float32).reshape(H, W, 3)
            want = orc.denoise(st, frame, bufs["BUF_RENDER_COLOR"].view(orc.COLOR_DTYPE), bufs["BUF_NORMAL"], bufs["BUF_DEPTH"], bufs["BUF_MOTION_VECTOR"], bufs["BUF_ALBEDO"])
            w = want["rgb"].astype(np.float32).reshape(H, W, 3)
            bad = ~np.isfinite(got).all(axis=2)
            badw = ~np.isfinite(w).all(axis=2)
            ys, xs = np.nonzero(bad)
            print(W, H, "frame", frame, "inputs finite", fin_in, "bad px", int(bad.sum()), "oracle bad", int(badw.sum()),
                  "first", list(zip(xs[:6].tolist(), ys[:6].tolist())), "x range", (int(xs.min()), int(xs.max())) if len(xs) else None,
                  "y range", (int(ys.min()), int(ys.max())) if len(ys) else None)
            for n in ("BUF_ACCUMULATION_COLOR", "BUF_HISTORY_COLOR"):
                a = rt.read_buffer(getattr(pkg, n))["rgb"].astype(np.float32)
                print("   ", n, "finite", bool(np.isfinite(a).all()))
            n16 = rt.read_buffer(pkg.BUF_NOISE_LEVEL_16).astype(np.float32)
            print("    noise16 finite", bool(np.isfinite(n16).all()), n16.shape)
