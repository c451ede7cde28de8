"""GPU tests of the row-sharded multi-GPU path.

* `test_row_bands_on_one_gpu`: N contexts on ONE device play the N ranks (the exchanges are device copies between their
  buffers): band path trace + G-buffer exchange + band-with-halo denoise + frame gather must reproduce, BIT FOR BIT, the
  frames a single whole-frame context renders — over several frames with a moving camera, so the temporal history that
  crosses band borders is exercised.
* `test_row_sharded_nccl`: the real thing over NCCL with one process per GPU (runs where >= 2 GPUs are visible).
"""
import importlib
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def dev_tensor(rt, pkg, which):
    import torch
    ptr, nbytes = rt.get_buffer(which)
    return torch.as_tensor(pkg.parallel._DevBytes(ptr, nbytes), device="cuda")


def exchange(rts, pkg, bufs, bands, W, rows_of=None):
    """what all_gather_rows / StripGather do over NCCL, as device copies between the contexts of one GPU.
    rows_of: per-rank row index arrays (interleaved strips); None = the contiguous bands"""
    import torch
    torch.cuda.synchronize()
    H = bands[-1][1]
    if rows_of is None:
        rows_of = [np.arange(b, e) for b, e in bands]
    for which, bpp in bufs:
        ts = [dev_tensor(rt, pkg, which)[: H * W * bpp].view(H, W * bpp) for rt in rts]
        for src, rows in enumerate(rows_of):
            idx = torch.from_numpy(np.asarray(rows, np.int64)).cuda()
            for dst in range(len(rts)):
                if dst != src:
                    ts[dst].index_copy_(0, idx, ts[src].index_select(0, idx))
    torch.cuda.synchronize()


@pytest.mark.parametrize("W,H,world,strip", [(256, 256, 2, 0), (200, 136, 3, 8), (320, 400, 4, 8), (256, 250, 2, 5)])
def test_row_bands_on_one_gpu(pkg, W, H, world, strip):
    pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C1"], seed=1)
    bands = pkg.row_bands(H, world)
    full = pkg.RayTracer(W, H).init()
    rts = [pkg.RayTracer(W, H).init() for _ in range(world)]
    try:
        for rt in [full] + rts:
            rt.set_mesh(pos, idx, nrm); rt.build_bvh()
        for r, rt in enumerate(rts):
            rt.set_row_range(*bands[r])
            if strip:
                rt.set_row_strips(strip, world, r)                      # path trace by interleaved strips, denoise by bands
        rows_of = pkg.parallel.strip_rows(H, strip, world) if strip else None
        for frame in range(1, 6):
            cam = pkg.scenes.make_camera((0.3 * frame, 30, -60 + 0.2 * frame), 0.01 * frame, -0.4 - 0.01 * frame, W, H)
            for rt in [full] + rts:
                rt.set_camera(cam, reset_history=(frame == 1)); rt.set_frame_number(frame)
                rt.pathtrace()
            exchange(rts, pkg, pkg.parallel.GBUFFERS, bands, W, rows_of)
            for which, _ in pkg.parallel.GBUFFERS:
                want = full.read_buffer(which)
                for rt in rts:
                    assert np.array_equal(rt.read_buffer(which).view(np.uint8), want.view(np.uint8)), f"frame {frame}: G-buffer {which}"
            rays = sum(rt.timings()["rays_traced"] for rt in rts)
            assert rays == full.timings()["rays_traced"]
            for rt in [full] + rts:
                rt.denoise()
            exchange(rts, pkg, pkg.parallel.DENOISED, bands, W)
            for which in (pkg.BUF_RENDER_COLOR, pkg.BUF_HISTORY_COLOR, pkg.BUF_ACCUMULATION_COLOR, pkg.BUF_HISTORY_DEPTH):
                want = full.read_buffer(which)
                for r, rt in enumerate(rts):
                    assert np.array_equal(rt.read_buffer(which).view(np.uint8), want.view(np.uint8)), f"frame {frame}: rank {r} buffer {which}"
    finally:
        for rt in [full] + rts:
            rt.cleanup()


def test_progressive_accumulation_by_bands(pkg):
    """BASELINE config 5 in miniature: 8 spp accumulated per band == accumulated on the whole frame, and == the mean of the frames."""
    W, H, spp = 192, 160, 8
    pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C1"], seed=1)
    cam = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, W, H)
    bands = pkg.row_bands(H, 2)
    full = pkg.RayTracer(W, H).init(); rts = [pkg.RayTracer(W, H).init() for _ in range(2)]
    try:
        ref_sum = np.zeros((W * H, 3), np.float64)
        for rt in [full] + rts:
            rt.set_mesh(pos, idx, nrm); rt.build_bvh(); rt.set_camera(cam, reset_history=True)
        for r, rt in enumerate(rts):
            rt.set_row_strips(8, 2, r)
        for s in range(spp):
            for rt in [full] + rts:
                rt.set_frame_number(1 + s); rt.pathtrace(); rt.accumulate(reset=(s == 0))
            ref_sum += full.read_buffer(pkg.BUF_RENDER_COLOR)["rgb"].astype(np.float64) * full.read_buffer(pkg.BUF_ALBEDO).astype(np.float64)[:, :3]
        exchange(rts, pkg, [(pkg.BUF_ACCUM_RADIANCE, 16)], bands, W, pkg.parallel.strip_rows(H, 8, 2))
        want = full.read_buffer(pkg.BUF_ACCUM_RADIANCE)
        assert (want[:, 3] == spp).all()
        np.testing.assert_allclose(want[:, :3], ref_sum, rtol=1e-5, atol=1e-6)
        for rt in rts:
            assert np.array_equal(rt.read_buffer(pkg.BUF_ACCUM_RADIANCE).view(np.uint8), want.view(np.uint8))
    finally:
        for rt in [full] + rts:
            rt.cleanup()


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _nccl_worker(rank, world, port, bvh_mode, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        pkg = importlib.import_module("real-time-ray-tracing_b200")
        W, H = 256, 256
        pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C1"], seed=1)
        rt = pkg.RayTracer(W, H, device=rank).init(); rt.set_mesh(pos, idx, nrm)
        sh = pkg.RowShardedRayTracer(rt, W, H, rank, world, bvh=bvh_mode)
        full = None
        if rank == 0:
            full = pkg.RayTracer(W, H, device=0).init(); full.set_mesh(pos, idx, nrm)
        for frame in range(1, 5):
            cam = pkg.scenes.make_camera((0.3 * frame, 30, -60 + 0.2 * frame), 0.01 * frame, -0.4, W, H)
            rt.set_camera(cam, reset_history=(frame == 1))
            sh.draw(frame)
            rt.sync()
            if rank == 0:
                full.set_camera(cam, reset_history=(frame == 1)); full.set_frame_number(frame)
                full.build_bvh(); full.pathtrace(); full.denoise()
                for which in (pkg.BUF_RENDER_COLOR, pkg.BUF_HISTORY_COLOR, pkg.BUF_ACCUMULATION_COLOR, pkg.BUF_HISTORY_DEPTH, pkg.BUF_NORMAL):
                    assert np.array_equal(rt.read_buffer(which).view(np.uint8), full.read_buffer(which).view(np.uint8)), f"frame {frame} buffer {which}"
        accs = {}
        for part in ("rows", "samples"):
            sh.render_progressive(4, first_frame=10, partition=part)
            rt.sync()
            accs[part] = rt.read_buffer(pkg.BUF_ACCUM_RADIANCE).copy()
            assert (accs[part][:, 3] == 4).all()
        np.testing.assert_allclose(accs["samples"][:, :3], accs["rows"][:, :3], rtol=1e-5, atol=1e-6)     # same samples, other summation order
        if rank == 0:
            open(os.path.join(out_dir, f"ok_{bvh_mode}"), "w").write("ok")
        rt.cleanup()
        if full is not None:
            full.cleanup()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("bvh_mode", ["rebuild", "broadcast"])
def test_row_sharded_nccl(bvh_mode, tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), 4)
    while 256 % world:
        world -= 1
    mp.spawn(_nccl_worker, args=(world, _free_port(), bvh_mode, str(tmp_path)), nprocs=world, join=True)
    assert (tmp_path / f"ok_{bvh_mode}").exists()
