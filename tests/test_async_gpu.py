"""Frames in flight (rtrt_update_vertices on the copy stream, rtrt_read_buffer_async, rtrt_join): the overlapped path must give
exactly the frames of the blocking path."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_pipelined_frames_equal_blocking_frames(pkg):
    W, H = 320, 192
    pos, nrm, idx = pkg.scenes.terrain(48, 48, seed=3)
    cam = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, W, H)
    rng = np.random.default_rng(1)
    anim = [(pos + rng.normal(scale=0.05, size=pos.shape).astype(np.float32)) for _ in range(4)]

    def run(pipelined):
        frames = []
        with pkg.RayTracer(W, H) as rt:
            rt.set_mesh(pos, idx, nrm); rt.set_camera(cam, reset_history=True)
            hp = [torch.from_numpy(a.reshape(-1).copy()).pin_memory() for a in anim]
            hn = torch.from_numpy(nrm.reshape(-1).copy()).pin_memory()
            out = [torch.empty(W * H * 8, dtype=torch.uint8).pin_memory() for _ in anim]
            for f, h in enumerate(hp):
                rt.set_frame_number(f + 1)
                rt.update_vertices(h.data_ptr(), hn.data_ptr())
                rt.build_bvh(); rt.pathtrace(); rt.denoise()
                if pipelined:
                    rt.read_buffer_async(pkg.BUF_RENDER_COLOR, out[f].data_ptr(), out[f].numel())      # host never blocks
                else:
                    rt.read_buffer_into(pkg.BUF_RENDER_COLOR, out[f].data_ptr(), out[f].numel())
            rt.join(); rt.sync()
            frames = [o.numpy().copy() for o in out]
            total = rt.timings()["rays_traced_total"]
        return frames, total

    a, ta = run(False)
    b, tb = run(True)
    assert ta == tb and ta > 4 * W * H
    for f, (x, y) in enumerate(zip(a, b)):
        np.testing.assert_array_equal(x, y, err_msg=f"frame {f + 1}")
    assert not np.array_equal(a[0], a[1])                       # the animation does change the frames


def test_async_read_argument_errors(pkg):
    with pkg.RayTracer(64, 64) as rt:
        buf = torch.empty(16, dtype=torch.uint8).pin_memory()
        with pytest.raises(pkg.RtrtError):
            rt.read_buffer_async(pkg.BUF_RENDER_COLOR, buf.data_ptr(), buf.numel())      # no frame yet / too small
        rt.join(); rt.sync()
