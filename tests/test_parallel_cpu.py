"""Host-side multi-GPU logic on the CPU (gloo, world_size 2): row bands, the in-place row all-gather (equal and ragged bands)
and the end-to-end sharding contract — two ranks path-trace their own row band with the CPU oracle, gather, and must hold
exactly the frame a single process renders."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_row_bands_cover_every_row_once(pkg):
    for H, world in [(1080, 1), (1080, 2), (1080, 8), (2160, 8), (45, 2), (7, 7), (1001, 4)]:
        bands = pkg.row_bands(H, world)
        assert len(bands) == world and bands[0][0] == 0 and bands[-1][1] == H
        assert all(b < e for b, e in bands) and all(bands[i][1] == bands[i + 1][0] for i in range(world - 1))
        assert max(e - b for b, e in bands) - min(e - b for b, e in bands) <= 1
    with pytest.raises(ValueError):
        pkg.row_bands(3, 4)


def test_assemble_rows(pkg):
    bands = pkg.row_bands(6, 2)
    a = np.arange(12).reshape(-1); b = -np.arange(12).reshape(-1)
    out = pkg.parallel.assemble_rows([a, b], bands, 2)
    assert out.tolist() == [0, 1, 2, 3, 4, 5, -6, -7, -8, -9, -10, -11]


def test_strip_rows_partition(pkg):
    for H, sh, world in [(1080, 8, 2), (1080, 8, 8), (2160, 8, 8), (45, 4, 2), (136, 8, 3), (17, 1, 4)]:
        per = pkg.parallel.strip_rows(H, sh, world)
        allr = np.sort(np.concatenate(per))
        assert np.array_equal(allr, np.arange(H))                        # every row exactly once
        for r, p in enumerate(per):
            assert ((p // sh) % world == r).all()


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, H, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pkg = importlib.import_module("real-time-ray-tracing_b200")
        from oracle import orc
        orc.build()
        W = 64
        pos, nrm, idx = pkg.scenes.terrain(32, 32, seed=2)
        ob = orc.build_bvh(pos, nrm, idx, threads=2)
        cam = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, W, H)
        sky = orc.sky_generate(orc.sun_dir())
        bands = pkg.row_bands(H, world)
        f = orc.pathtrace(ob, cam, W, H, 3, sky, threads=2, rows=bands[rank])           # this rank's band only
        planes = {"color": f.color, "normal": f.normal, "albedo": f.albedo, "depth": f.depth, "motion": f.motion, "mask": f.mask}
        gathered = {}
        for name, a in planes.items():
            row_bytes = a.nbytes // H
            t = torch.from_numpy(a.view(np.uint8).reshape(-1).copy())
            b, e = bands[rank]
            other = torch.ones_like(t) * 0xAB                                          # rows this rank does not own: poison
            other[b * row_bytes: e * row_bytes] = t[b * row_bytes: e * row_bytes]
            pkg.parallel.all_gather_rows(other, row_bytes, bands, rank)
            gathered[name] = other.numpy().view(a.dtype).reshape(a.shape)
        # the interleaved-strip gather on the same planes: keep only the own strips, poison the rest, gather
        sg = pkg.parallel.StripGather(H, 4, world, rank, "cpu")
        full_planes = {k: v.copy() for k, v in gathered.items()}
        for name, a in full_planes.items():
            row_bytes = a.nbytes // H
            t = torch.from_numpy(a.view(np.uint8).reshape(-1).copy())
            keep = np.zeros(H, bool); keep[pkg.parallel.strip_rows(H, 4, world)[rank]] = True
            t.view(H, row_bytes)[torch.from_numpy(~keep)] = 0xCD
            sg.gather(t, row_bytes)
            assert np.array_equal(t.numpy(), a.view(np.uint8).reshape(-1)), f"{name}: strip gather"
        if rank == 0:
            full = orc.pathtrace(ob, cam, W, H, 3, sky, threads=2)
            for name, a in gathered.items():
                want = getattr(full, name)
                assert np.array_equal(a.view(np.uint8), want.view(np.uint8)), f"{name}: gathered bands differ from the single-process frame"
            open(os.path.join(out_dir, f"ok_{H}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("H", [48, 45])        # equal bands (all_gather_into_tensor) and ragged bands (per-rank broadcast)
def test_two_ranks_render_the_single_process_frame(H, tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, H, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / f"ok_{H}").exists()
