#!/usr/bin/env python
"""Per-kernel times of one path-traced frame (CUDA events are too coarse: run under `ncu --metrics gpu__time_duration.sum`)."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
pkg = importlib.import_module("real-time-ray-tracing_b200")
W, H = 1920, 1080
pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES[os.environ.get("SCENE", "C2ref")], seed=1)
cam = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, W, H)
rt = pkg.RayTracer(W, H).init(); rt.set_mesh(pos, idx, nrm); rt.set_camera(cam, reset_history=True)
for k, v in [a.split("=") for a in sys.argv[1:]]:
    rt.set_option(k, int(v))
rt.build_bvh()
for f in range(2, 5):
    rt.set_frame_number(f); rt.pathtrace()
rt.sync()
print(rt.timings()["rays_traced"])
if os.environ.get("DIAG"):
    rt.set_option("count_visits", 1); rt.set_frame_number(2); rt.pathtrace(); rt.sync()
    c = rt.read_buffer(pkg.BUF_WF_COUNTERS)
    for r in range(4):
        n = int(c[r])
        print(f"round {r+1}: rays {n} hits {int(c[16+r])} misses {int(c[24+r])} max_it {int(c[40+r])} mean_it {int(c[48+r]) / max(n,1):.1f}")
