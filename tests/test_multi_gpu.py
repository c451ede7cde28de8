"""GPU tests of the C-level multi-GPU frame (csrc/multi.cu: rtrt_render_frame_multi and friends).

* in-process transport (rtrt_multi_link_local): N contexts on ONE device, each driven by its own host thread, play the N ranks with the
  same exchange plans, pack / unpack kernels and denoiser halo hook as over NCCL.  Strips for the path tracer + bands with a halo exchange
  before every filter pass must reproduce a single context's frames BIT FOR BIT (denoised colour, accumulation, history) on every rank;
* sample partition: the mean of the ranks' frames; progressive accumulation by samples vs a single context (summation order differs);
* the same over real NCCL, one process per GPU (skipped on a single-GPU box)."""
import importlib
import os
import socket
import sys
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run_ranks(fns):
    """one host thread per rank (ctypes releases the GIL inside the library); re-raises the first failure"""
    errs = [None] * len(fns)

    def wrap(i):
        try:
            fns[i]()
        except BaseException as e:          # noqa: BLE001
            errs[i] = e
    ts = [threading.Thread(target=wrap, args=(i,)) for i in range(len(fns))]
    for t in ts: t.start()
    for t in ts: t.join(timeout=120)
    assert not any(t.is_alive() for t in ts), "a rank hangs in an exchange"
    for e in errs:
        if e is not None:
            raise e


def _cameras(pkg, W, H, n):
    return [pkg.scenes.make_camera((0.3 * f, 30, -60 + 0.2 * f), 0.01 * f, -0.4, W, H) for f in range(1, n + 1)]


@pytest.mark.parametrize("W,H,world,passes", [(256, 256, 2, (1, 1, 1, 1)), (320, 400, 3, (1, 1, 1, 1)), (256, 256, 4, (1, 1, 2, 1)), (200, 136, 2, (1, 0, 1, 1))])
def test_rows_mode_in_process_is_bit_identical_to_one_context(pkg, W, H, world, passes):
    pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C1"], seed=1)
    full = pkg.RayTracer(W, H).init(); full.set_mesh(pos, idx, nrm)
    rts = [pkg.RayTracer(W, H).init() for _ in range(world)]
    try:
        for rt in rts + [full]:
            rt.set_denoise_params(None, pkg.PassSettings(*passes))
        for rt in rts:
            rt.set_mesh(pos, idx, nrm)
        pkg.parallel.link_local(rts)
        lib = rts[0]._lib
        for frame, cam in enumerate(_cameras(pkg, W, H, 4), start=1):
            full.set_camera(cam, reset_history=(frame == 1)); full.set_frame_number(frame)
            full.build_bvh(); full.pathtrace(); full.denoise()

            def rank_fn(rt):
                def f():
                    rt.set_camera(cam, reset_history=(frame == 1)); rt.set_frame_number(frame)
                    rt._check(lib.rtrt_render_frame_multi(rt._h)); rt.sync()
                return f
            _run_ranks([rank_fn(rt) for rt in rts])
            want = {b: full.read_buffer(b).view(np.uint8) for b in (pkg.BUF_RENDER_COLOR, pkg.BUF_HISTORY_COLOR, pkg.BUF_ACCUMULATION_COLOR)}
            for r, rt in enumerate(rts):
                for b, w in want.items():
                    if b == pkg.BUF_ACCUMULATION_COLOR and not passes[0]:
                        continue
                    if b == pkg.BUF_HISTORY_COLOR and not passes[3]:
                        continue
                    got = rt.read_buffer(b).view(np.uint8)
                    bad = np.nonzero((got != w).reshape(H, -1).any(axis=1))[0]
                    assert bad.size == 0, f"frame {frame} rank {r} buffer {b}: rows {bad[:10]} differ from the single-context frame"
        import ctypes as C
        x, n = C.c_uint64(), C.c_uint64()
        rts[0]._check(lib.rtrt_multi_get_band(rts[0]._h, None, None, C.byref(x), C.byref(n)))
        assert x.value >= 4 * 3 and n.value > 0                   # G, halos, F every frame
    finally:
        for rt in rts + [full]:
            rt.cleanup()


def test_samples_mode_and_progressive_in_process(pkg):
    W, H, world = 192, 160, 3
    pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C1"], seed=1)
    cam = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, W, H)
    full = pkg.RayTracer(W, H).init(); full.set_mesh(pos, idx, nrm); full.set_camera(cam, reset_history=True)
    rts = [pkg.RayTracer(W, H).init() for _ in range(world)]
    try:
        for rt in rts:
            rt.set_mesh(pos, idx, nrm); rt.set_camera(cam, reset_history=True)
        pkg.parallel.link_local(rts)
        lib = rts[0]._lib
        for rt in rts:
            rt._check(lib.rtrt_multi_set_mode(rt._h, pkg._lib.MULTI_SAMPLES, 8))
        # frame 1 of the group = samples 1, 2, 3 of the single context (first frames: no temporal history yet on any of them)
        singles = []
        for s in range(world):
            one = pkg.RayTracer(W, H).init(); one.set_mesh(pos, idx, nrm); one.set_camera(cam, reset_history=True)
            one.set_frame_number(1 + s); one.build_bvh(); one.pathtrace(); one.denoise()
            singles.append(one.read_buffer(pkg.BUF_RENDER_COLOR)); one.cleanup()

        def rank_fn(rt):
            def f():
                rt.set_frame_number(1); rt._check(lib.rtrt_render_frame_multi(rt._h)); rt.sync()
            return f
        _run_ranks([rank_fn(rt) for rt in rts])
        mean = np.mean([s["rgb"].astype(np.float32) for s in singles], axis=0)
        for rt in rts:
            got = rt.read_buffer(pkg.BUF_RENDER_COLOR)
            np.testing.assert_allclose(got["rgb"].astype(np.float32), mean, rtol=2e-3, atol=2e-3)
            assert np.array_equal(got["mask"], singles[rts.index(rt)]["mask"])          # the mask lane is the rank's own, not averaged bits
        # progressive: 7 samples over 3 ranks vs one context
        for s in range(7):
            full.set_frame_number(10 + s); full.build_bvh() if s == 0 else None; full.pathtrace(); full.accumulate(reset=(s == 0))
        want = full.read_buffer(pkg.BUF_ACCUM_RADIANCE)

        def prog_fn(rt):
            def f():
                rt._check(lib.rtrt_render_progressive_multi(rt._h, 7, 10)); rt.sync()
            return f
        _run_ranks([prog_fn(rt) for rt in rts])
        for rt in rts:
            got = rt.read_buffer(pkg.BUF_ACCUM_RADIANCE)
            assert (got[:, 3] == 7).all()
            np.testing.assert_allclose(got[:, :3], want[:, :3], rtol=1e-5, atol=1e-6)
    finally:
        for rt in rts + [full]:
            rt.cleanup()


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _nccl_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)            # only hands out the NCCL id; the data plane is the library's own NCCL
    try:
        pkg = importlib.import_module("real-time-ray-tracing_b200")
        W, H = 256, 256
        pos, nrm, idx = pkg.scenes.terrain(*pkg.scenes.TERRAIN_SIZES["C1"], seed=1)
        rt = pkg.RayTracer(W, H, device=rank).init(); rt.set_mesh(pos, idx, nrm)
        mg = pkg.parallel.MultiGpuRayTracer.bootstrap(rt, mode="rows")
        full = pkg.RayTracer(W, H, device=rank).init(); full.set_mesh(pos, idx, nrm)
        # rtrt_create_multi = rtrt_create + rtrt_multi_init in one call (a second communicator on the same GPUs); a bad rank is an error, not a hang
        import ctypes as C
        box = [pkg.parallel.MultiGpuRayTracer.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        h = C.c_void_p()
        cfg = pkg.config_defaults(); cfg.width, cfg.height = W, H
        buf = (C.c_ubyte * 128).from_buffer_copy(box[0])
        lib = pkg._lib.load()
        assert lib.rtrt_create_multi(C.byref(cfg), rank, buf, world, world, C.byref(h)) == -1 and not h       # rank out of range -> RTRT_ERR_ARG
        assert lib.rtrt_create_multi(C.byref(cfg), rank, buf, rank, world, C.byref(h)) == 0 and h
        b, e = C.c_int32(), C.c_int32()
        assert lib.rtrt_multi_get_band(h, C.byref(b), C.byref(e), None, None) == 0 and (b.value, e.value) == pkg.parallel.band_bounds(H, world)[rank]
        assert lib.rtrt_destroy(h) == 0
        for frame in range(1, 5):
            cam = pkg.scenes.make_camera((0.3 * frame, 30, -60 + 0.2 * frame), 0.01 * frame, -0.4, W, H)
            rt.set_camera(cam, reset_history=(frame == 1))
            mg.draw(frame); rt.sync()
            full.set_camera(cam, reset_history=(frame == 1)); full.set_frame_number(frame)
            full.build_bvh(); full.pathtrace(); full.denoise()
            for which in (pkg.BUF_RENDER_COLOR, pkg.BUF_HISTORY_COLOR, pkg.BUF_ACCUMULATION_COLOR):
                assert np.array_equal(rt.read_buffer(which).view(np.uint8), full.read_buffer(which).view(np.uint8)), f"rank {rank} frame {frame} buffer {which}"
        # vertex broadcast: rank 0 uploads, everyone rebuilds from the broadcast buffers
        import ctypes as C
        p2 = np.ascontiguousarray(pos + np.float32([0, 0.25, 0])); n2 = np.ascontiguousarray(nrm)
        mg.update_vertices(p2.ctypes.data_as(C.c_void_p) if rank == 0 else None, n2.ctypes.data_as(C.c_void_p) if rank == 0 else None)
        mg.draw(5); rt.sync()
        full.set_mesh(p2, idx, n2); full.set_frame_number(5); full.build_bvh(); full.pathtrace()
        assert np.array_equal(rt.read_buffer(pkg.BUF_TRIANGLES).view(np.uint8), full.read_buffer(pkg.BUF_TRIANGLES).view(np.uint8))
        mg.render_progressive(6, first_frame=20); rt.sync()
        acc = rt.read_buffer(pkg.BUF_ACCUM_RADIANCE)
        assert (acc[:, 3] == 6).all()
        mg.set_mode("samples")
        rt.set_frame_number(30); mg.draw(); rt.sync()
        assert np.isfinite(rt.read_buffer(pkg.BUF_RENDER_COLOR)["rgb"].astype(np.float32)).all()
        open(os.path.join(out_dir, f"ok_{rank}"), "w").write("ok")
        rt.cleanup(); full.cleanup()
    finally:
        dist.destroy_process_group()


def test_rows_mode_over_nccl(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), 4)
    mp.spawn(_nccl_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok_{r}").exists() for r in range(world))
