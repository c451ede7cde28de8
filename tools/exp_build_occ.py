#!/usr/bin/env python
"""Build-kernel timing + achieved occupancy (experiment helper)."""
import importlib, os, sys, ctypes
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
pkg = importlib.import_module("real-time-ray-tracing_b200")
import torch
for scene in ("C2ref", "C3"):
    if scene not in pkg.scenes.TERRAIN_SIZES: continue
    pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES[scene], seed=1)
    rt = pkg.RayTracer(64, 64).init(); rt.set_mesh(pos, idx, nrm)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ts = []
    for i in range(12):
        flush.zero_(); torch.cuda.synchronize()
        rt.build_bvh(); rt.sync()
        ts.append(rt.timings()["build_bvh_ms"])
    print(scene, idx.shape[0], "build ms median", float(np.median(ts[2:])), "min", min(ts[2:]))
    rt.cleanup()
