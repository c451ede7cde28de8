"""Wide-layout statistics of the bench scene: nodes written, children per node, bytes."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
pkg = importlib.import_module("real-time-ray-tracing_b200")
pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES[sys.argv[1] if len(sys.argv) > 1 else "C2ref"], seed=1)
with pkg.RayTracer(64, 64) as rt:
    rt.set_option("wide_bvh", 1); rt.set_mesh(pos, idx, nrm); rt.build_bvh(); rt.sync()
    stat = rt.read_buffer(pkg.BUF_WIDE_STAT)
    raw = rt.read_buffer(pkg.BUF_WIDE_NODES)
    nb = rt.counts()[1]
meta = raw[:, 24:32]
used = np.zeros(raw.shape[0], bool)
# nodes in use: per batch region the first k nodes are allocated breadth-first; count nodes with any non-empty slot
nonempty = (meta != 0).sum(axis=1)
inuse = nonempty > 0
print("triangles", idx.shape[0], "batches", nb, "stat", stat.tolist(), "nodes with children", int(inuse.sum()), "children per node %.2f" % nonempty[inuse].mean(),
      "inner slots %.2f leaf slots %.2f" % (((meta & 0x18) == 0x18)[inuse].sum(axis=1).mean(), (((meta != 0) & ((meta & 0x18) != 0x18))[inuse]).sum(axis=1).mean()),
      "node MB %.1f" % (inuse.sum() * 128 / 1e6))
