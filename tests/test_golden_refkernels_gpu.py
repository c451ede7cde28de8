"""GPU parity tests of the PRODUCT (CUDA kernels through the C ABI) against the committed golden vectors made by the reference's own kernels
(tests/golden/refkernels_c1.npz, tools/make_golden_refkernels.py) — no oracle, no oracle/_ref at run time.

Stage 1 and the per-ray hit records: bit for bit.  Path-traced G-buffers of frame 1: the tolerances of tests/test_refcuda_frame_gpu.py
(the golden frames sampled the CPU oracle's sky maps, the product samples the maps its own kernels generate: <= 2e-4 of the map maximum
apart, tests/test_refcuda_frame_gpu.py::test_sky_maps_match_reference_kernels).
Not compared here: the denoised golden frames.  They are the reference denoiser's output on the REFERENCE's path-traced frame; the product's
1-spp frame differs from it in a few pixels whose discrete path differs, and the filters spread those (measured: rel-RMSE 2.7e-2 at
48x48, against the 2e-2 bound that holds on identical G-buffers).  The denoiser is compared on identical G-buffers, live against the
reference kernels, in tests/test_refcuda_frame_gpu.py and tests/test_bench_config_gpu.py; the CPU oracle is held to the golden denoised
frames in tests/test_golden_refkernels_cpu.py."""
import json
import os

import numpy as np
import pytest

from helpers import GOLDEN_H, GOLDEN_W, assert_hits_equal, golden_refkernels_inputs, golden_stage1_digests, read_all_stage1

pytestmark = pytest.mark.gpu
GOLD_PATH = os.path.join(os.path.dirname(__file__), "golden", "refkernels_c1.npz")


def rel_rmse(a, b):
    a = a.astype(np.float64); b = b.astype(np.float64)
    return float(np.sqrt(np.mean((a - b) ** 2)) / max(np.sqrt(np.mean(b ** 2)), 1e-12))


@pytest.fixture(scope="module")
def gold():
    g = dict(np.load(GOLD_PATH))
    g["meta"] = json.loads(str(g["meta"]))
    return g


class _Stage1:
    pass


def test_product_matches_reference_kernel_golden_vectors(pkg, gold):
    wide = 0                                   # the reference's layout and loop: every record bit for bit
    inp = golden_refkernels_inputs(pkg)
    W, H = GOLDEN_W, GOLDEN_H
    with pkg.RayTracer(W, H) as rt:
        rt.set_option("wide_bvh", wide)
        rt.set_mesh(inp["pos"], inp["idx"], inp["nrm"])
        rt.build_bvh()
        # ---- stage 1: every observation point, bit for bit ----
        bufs = read_all_stage1(rt, pkg)
        s = _Stage1()
        s.num_batches = (inp["idx"].shape[0] + 1023) // 1024
        s.triangles, s.morton, s.reorder_idx, s.nodes = bufs[pkg.BUF_TRIANGLES], bufs[pkg.BUF_MORTON], bufs[pkg.BUF_REORDER_IDX], bufs[pkg.BUF_BVH_NODES]
        s.tlas_aabbs, s.tlas_morton, s.tlas_reorder_idx, s.tlas_nodes = (bufs[pkg.BUF_TLAS_AABBS], bufs[pkg.BUF_TLAS_MORTON], bufs[pkg.BUF_TLAS_REORDER_IDX],
                                                                          bufs[pkg.BUF_TLAS_BVH_NODES])
        got, want = golden_stage1_digests(s), gold["meta"]["digests"]
        bad = [k for k in sorted(want) if got[k] != want[k]]
        assert not bad, f"product differs from the reference kernels at: {bad}"
        # ---- traversal: (tri, t, u, v) of every ray, bit for bit ----
        hits = np.zeros(gold["hit_tri"].shape[0], np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("tri", "<i4")]))
        for f in ("t", "u", "v", "tri"):
            hits[f] = gold["hit_" + f]
        assert_hits_equal(rt.traverse(inp["rays"]), hits)
        # ---- path trace, frame 1 ----
        for frame, cam in inp["frames"][:1]:
            rt.set_camera(cam, reset_history=(frame == 1)); rt.set_frame_number(frame)
            rt.pathtrace()
            color, depth = rt.read_buffer(pkg.BUF_RENDER_COLOR), rt.read_buffer(pkg.BUF_DEPTH).astype(np.float32)
            wcolor = np.ascontiguousarray(gold[f"f{frame}_color"]).view(color.dtype).reshape(-1)
            wd = gold[f"f{frame}_depth"].view("<f2").astype(np.float32)
            mask_ok = color["mask"] == wcolor["mask"]
            same_hit = np.isinf(depth) == np.isinf(wd)
            assert (~mask_ok).sum() <= 3 and (~same_hit).sum() <= 3, frame
            hit = np.isfinite(wd) & same_hit
            assert rel_rmse(depth[hit], wd[hit]) <= 1e-3, frame
            rgb, wrgb = color["rgb"].astype(np.float32), wcolor["rgb"].astype(np.float32)
            d = np.abs(rgb - wrgb).max(axis=1) / (np.abs(wrgb).max(axis=1) + 1e-3)
            agree = (d < 0.05) & mask_ok & same_hit
            e_rgb = rel_rmse(rgb[agree], wrgb[agree])
            e_alb = rel_rmse(rt.read_buffer(pkg.BUF_ALBEDO).astype(np.float32)[agree, :3], gold[f"f{frame}_albedo"].view("<f2").astype(np.float32)[agree, :3])
            e_nrm = rel_rmse(rt.read_buffer(pkg.BUF_NORMAL).astype(np.float32)[agree, :3], gold[f"f{frame}_normal"].view("<f2").astype(np.float32)[agree, :3])
            e_mot = rel_rmse(rt.read_buffer(pkg.BUF_MOTION_VECTOR).astype(np.float32)[agree], gold[f"f{frame}_motion"].view("<f2").astype(np.float32)[agree])
            print(f"golden wide={wide} frame {frame}: agree {agree.mean():.4f} exact {(d == 0).mean():.4f} rgb {e_rgb:.2e} albedo {e_alb:.2e} "
                  f"normal {e_nrm:.2e} motion {e_mot:.2e}")
            assert agree.mean() >= 0.995, f"frame {frame}: only {agree.mean():.4f} of the pixels follow the reference kernel's path"
            assert e_rgb <= 1e-3 and e_alb <= 1e-3 and e_nrm <= 2e-3 and e_mot <= 1e-3, frame
