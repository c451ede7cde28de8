#!/usr/bin/env python
"""A/B of the denoiser on the bench frame: median denoise stage time (CUDA events of the library, L2 flushed) and a SHA-1 of the
denoised frames (1080p bench scene, and a 250x130 frame whose tiles are mostly border tiles), so that builds which must be
bit-identical can be told apart from builds that are not.
usage: [RTRT_EXPERIMENT=1 RTRT_LIB=...] python tools/exp_dn_ab.py [passes=a,b,c,d]"""
import hashlib, importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
pkg = importlib.import_module("real-time-ray-tracing_b200")
passes = None
for a in sys.argv[1:]:
    if a.startswith("passes="): passes = tuple(int(v) for v in a.split("=")[1].split(","))
pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C2ref"], seed=1)
out = []
for (W, H, frames) in ((1920, 1080, 12), (250, 130, 4)):
    cam = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, W, H)
    rt = pkg.RayTracer(W, H).init(); rt.set_mesh(pos, idx, nrm); rt.set_camera(cam, reset_history=True)
    rt.set_option("wide_bvh", 1)
    if passes: rt.set_denoise_params(None, pkg.PassSettings(*passes))
    stream = torch.cuda.ExternalStream(rt.stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    h = hashlib.sha1(); ms = []
    warm = []
    for i in range(2 * frames):
        rt.set_frame_number(1 + i); rt.build_bvh(); rt.pathtrace()
        if i < frames:
            with torch.cuda.stream(stream): flush.zero_()
        rt.denoise()
        (ms if i < frames else warm).append(rt.timings()["denoise_ms"])
        if i < 4: h.update(rt.read_buffer(pkg.BUF_RENDER_COLOR).tobytes())
    out.append("%dx%d denoise %.4f ms (L2 flushed) %.4f ms (as in the frame) sha1 %s" % (W, H, float(np.median(ms[2:])), float(np.median(warm)), h.hexdigest()[:16]))
    rt.cleanup()
print(os.path.basename(os.environ.get("RTRT_LIB", "in-tree")), " | ".join(out), flush=True)
