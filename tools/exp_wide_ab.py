#!/usr/bin/env python
"""A/B of the path-trace kernels on the bench scene: median per-kernel CUDA-event times (library option kernel_timing), L2 flushed.
usage: [RTRT_EXPERIMENT=1 RTRT_LIB=...] python tools/exp_wide_ab.py opt=val ..."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
pkg = importlib.import_module("real-time-ray-tracing_b200")
W, H = 1920, 1080
pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C2ref"], seed=1)
cam = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, W, H)
rt = pkg.RayTracer(W, H).init(); rt.set_mesh(pos, idx, nrm); rt.set_camera(cam, reset_history=True)
for k, v in [a.split("=") for a in sys.argv[1:]]:
    rt.set_option(k, int(v))
rt.set_option("kernel_timing", 1)
stream = torch.cuda.ExternalStream(rt.stream)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
rows, builds, wides = [], [], []
for i in range(14):
    with torch.cuda.stream(stream): flush.zero_()
    rt.build_bvh(); rt.set_frame_number(2 + i); rt.pathtrace()
    t = rt.timings()
    rows.append(t["pt_kernel_ms"] + [t["pathtrace_ms"]]); builds.append(t["build_bvh_ms"]); wides.append(t["wide_build_ms"])
m = np.median(np.asarray(rows[2:]), axis=0)
rt.set_option("count_visits", 1); rt.set_frame_number(2); rt.pathtrace(); t = rt.timings()
print(os.path.basename(os.environ.get("RTRT_LIB", "in-tree")), " ".join(sys.argv[1:]), "| build %.3f + collapse %.3f | primary %.3f logic1 %.3f trace1 %.3f logic %.3f trace2 %.3f rest %.3f | pathtrace %.3f ms | r1 visits/ray %.1f tris/ray %.1f" %
      (float(np.median(builds[2:])), float(np.median(wides[2:])), m[0], m[1], m[2], m[3], m[4], m[5], m[6], t["round_node_visits"][0] / max(t["round_rays"][0], 1), t["round_leaf_tests"][0] / max(t["round_rays"][0], 1)), flush=True)
rt.cleanup()
