#!/usr/bin/env python
"""C3 with shuffled input: device Morton pre-sort + build (for an ncu launch list / timing of the sort passes)."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
pkg = importlib.import_module("real-time-ray-tracing_b200")
pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C3"], seed=1, order="shuffle")
rt = pkg.RayTracer(64, 64).init(); rt.set_option("global_sort", 1); rt.set_mesh(pos, idx, nrm)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ts = []
for i in range(int(os.environ.get("N", 8))):
    flush.zero_(); torch.cuda.synchronize()
    rt.build_bvh(); rt.sync()
    ts.append(rt.timings()["build_bvh_ms"])
print("C3 shuffled", idx.shape[0], "build ms median", float(np.median(ts[2:])), "min", min(ts[2:]))
rt.cleanup()
