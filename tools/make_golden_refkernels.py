#!/usr/bin/env python
"""Golden vectors from the REFERENCE'S OWN CUDA kernels (oracle/_ref/librefcuda.so, built from /root/reference/src by
oracle/refcuda/build_ref.py) for the CPU suite: tests/test_golden_refkernels_cpu.py checks the CPU oracle against them without a GPU.

    python tools/make_golden_refkernels.py [--out tests/golden/refkernels_c1.npz]        (needs a GPU: run under gpurun)
    python tools/make_golden_refkernels.py --backend oracle --out /tmp/self.npz            (dry run of the script on the CPU oracle)

Stored: SHA-1 digests of every stage-1 observation point (UpdateSceneGeometry / RadixSort / BuildLBVH / UpdateTLAS), the (t, u, v, tri)
records of TraverseBvh for 4 072 rays, and for three 48x48 frames the G-buffers PathTrace wrote and the colour image after
TemporalSpatialDenoising.  Inputs are not stored: tests/helpers.py::golden_refkernels_inputs regenerates them.  The sky maps the
reference kernels sample are the CPU oracle's (orc.sky_generate; the reference's Sky / SkySun / Scan kernels are compared with them
on the GPU in tests/test_refcuda_frame_gpu.py), because 2 MB of maps do not belong in a fixture."""
import argparse, importlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
pkg = importlib.import_module("real-time-ray-tracing_b200")
from oracle import orc
import helpers

ap = argparse.ArgumentParser()
ap.add_argument("--out", default=os.path.join(ROOT, "tests", "golden", "refkernels_c1.npz"))
ap.add_argument("--backend", default="reference", choices=["reference", "oracle"])
args = ap.parse_args()
W, H = helpers.GOLDEN_W, helpers.GOLDEN_H
inp = helpers.golden_refkernels_inputs(pkg)
pos, nrm, idx, rays = inp["pos"], inp["nrm"], inp["idx"], inp["rays"]
N = idx.shape[0]
sky = orc.sky_generate(orc.sun_dir())
out = {}


def put_frame(k, g, den):
    out[f"f{k}_color"] = np.ascontiguousarray(g["color"]).view(np.uint16).reshape(-1, 4)
    out[f"f{k}_normal"] = np.ascontiguousarray(g["normal"]).view(np.uint16).reshape(-1, 4)
    out[f"f{k}_albedo"] = np.ascontiguousarray(g["albedo"]).view(np.uint16).reshape(-1, 4)
    out[f"f{k}_depth"] = np.ascontiguousarray(g["depth"]).view(np.uint16).reshape(-1)
    out[f"f{k}_motion"] = np.ascontiguousarray(g["motion"]).view(np.uint16).reshape(-1, 2)
    out[f"f{k}_denoised"] = np.ascontiguousarray(den).view(np.uint16).reshape(-1, 4)


if args.backend == "reference":
    from oracle.refcuda import ref

    class Stage1:
        pass
    rb = ref.build_bvh(pos, nrm, idx)
    s = Stage1()
    s.num_batches = rb.num_batches
    s.triangles = rb.host("triangles", orc.TRIANGLE_DTYPE); s.morton = rb.host("morton", np.uint32); s.reorder_idx = rb.host("reorderIdx", np.uint32)
    s.nodes = rb.host("nodes", orc.BVHNODE_DTYPE); s.tlas_aabbs = rb.host("tlasAabbs", orc.AABB_DTYPE); s.tlas_morton = rb.host("tlasMorton", np.uint32)
    s.tlas_reorder_idx = rb.host("tlasReorderIdx", np.uint32); s.tlas_nodes = rb.host("tlasNodes", orc.BVHNODE_DTYPE)
    digests = helpers.golden_stage1_digests(s)
    hits, _ = ref.traverse(rb, rays)
    bn, _ = orc.load_tables()
    mats = orc.default_materials()
    midx = np.full(rb.num_batches * 1024, 3, np.int32)
    with ref.RefFrame(W, H) as rf:
        rf.set_tables(bn); rf.set_scene(rb); rf.set_materials(mats, midx)
        rf.set_texture(0, orc.mip_chain(orc.DEFAULT_TEX[0])); rf.set_texture(1, orc.mip_chain(orc.DEFAULT_TEX[1]))
        rf.set_sky(sky)
        hist = None
        for frame, cam in inp["frames"]:
            rf.pathtrace(cam, frame, history_camera=hist)
            g = {"color": rf.read(ref.RB_COLOR), "normal": rf.read(ref.RB_NORMAL), "albedo": rf.read(ref.RB_ALBEDO),
                 "depth": rf.read(ref.RB_DEPTH), "motion": rf.read(ref.RB_MOTION)}
            rf.denoise(frame)
            put_frame(frame, g, rf.read(ref.RB_COLOR))
            hist = cam
    made_by = "oracle/_ref/librefcuda.so (--fmad=false): refcuda_build_bvh, refcuda_traverse, refcuda_frame_pathtrace, refcuda_frame_denoise"
else:
    from test_refcuda_frame_gpu import oracle_gbuffers
    ob = orc.build_bvh(pos, nrm, idx)
    digests = helpers.golden_stage1_digests(ob)
    hits = orc.traverse(ob, rays)
    st = orc.DenoiseState(W, H)
    hist = None
    for frame, cam in inp["frames"]:
        g = oracle_gbuffers(orc.pathtrace(ob, cam, W, H, frame, sky, history_camera=hist), orc)
        den = orc.denoise(st, frame, g["color"].view(orc.COLOR_DTYPE), g["normal"], g["depth"], g["motion"], g["albedo"])
        put_frame(frame, g, den)
        hist = cam
    made_by = "CPU oracle (dry run of the script; NOT a golden vector)"

out["hit_t"] = np.ascontiguousarray(hits["t"]); out["hit_u"] = np.ascontiguousarray(hits["u"]); out["hit_v"] = np.ascontiguousarray(hits["v"])
out["hit_tri"] = np.ascontiguousarray(hits["tri"])
out["meta"] = np.array(json.dumps({"made_by": made_by, "digests": digests, "num_tris": int(N), "num_rays": int(rays.shape[0]), "width": W, "height": H,
                                   "hit_rate": float((hits["tri"] >= 0).mean())}))
os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
np.savez_compressed(args.out, **out)
print("wrote", args.out, os.path.getsize(args.out), "bytes;", made_by, "; hit rate %.3f" % float((hits["tri"] >= 0).mean()))
