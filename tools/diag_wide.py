"""Diagnostic: rays on which the wide layout differs from the uncapped reference-layout search (C1 scene of tests/test_wide_bvh_gpu.py)."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
pkg = importlib.import_module("real-time-ray-tracing_b200")
from test_wide_bvh_gpu import _random_rays, _bounce_rays

pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C1"], seed=1)
cam = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, 256, 256)
rays = np.concatenate([pkg.scenes.primary_rays(cam, 256, 256), _random_rays(pkg, 100000, 7, (-60, -10, -60), (60, 20, 60)), _bounce_rays(pkg, pos, idx, 200000, 9)])
rays["dir"][65536:65836] = np.tile(np.eye(3, dtype=np.float32), (100, 1))
rays["dir"][65836:66136] = -np.tile(np.eye(3, dtype=np.float32), (100, 1))
with pkg.RayTracer(64, 64) as rt:
    rt.set_mesh(pos, idx, nrm); rt.build_bvh()
    rt.set_option("stack_size", 64); rt.set_option("max_loop", 1 << 20)
    deep = rt.traverse(rays)
    rt.set_option("wide_bvh", 1); rt.build_bvh()
    wide = rt.traverse(rays)
diff = (wide["tri"] != deep["tri"]) | (wide["t"].view(np.uint32) != deep["t"].view(np.uint32))
ids = np.nonzero(diff)[0]
print(os.environ.get("RTRT_LIB", "in-tree"), "differ:", ids.shape[0], "wide missed closer:", int((wide["t"][ids] > deep["t"][ids]).sum()), "wide found closer:", int((wide["t"][ids] < deep["t"][ids]).sum()))
print("index ranges:", {"primary": int((ids < 65536).sum()), "axis": int(((ids >= 65536) & (ids < 66136)).sum()), "random": int(((ids >= 66136) & (ids < 165536)).sum()), "bounce": int((ids >= 165536).sum())})
for i in ids[:12]:
    print(i, rays["orig"][i], rays["dir"][i], "wide", wide["tri"][i], wide["t"][i], "deep", deep["tri"][i], deep["t"][i])
