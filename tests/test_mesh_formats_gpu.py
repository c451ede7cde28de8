"""SURVEY 8f N4 on the GPU: OBJ / .bin ingest into the build, the `.bin` writer with the device 60-bit Morton pre-sort
(tool/meshProcessor.cpp:146-215) against the oracle restatement and against the digest of the reference's own test.bin."""
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import mesh_formats as mf
from test_mesh_formats_cpu import OBJ_TEXT, write_obj

pytestmark = pytest.mark.gpu
GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "mesh_formats.json")))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, np.float32).tobytes()).hexdigest()


def soup(pos):
    n = pos.shape[0]
    return pos.reshape(-1, 3).copy(), np.arange(3 * n, dtype=np.uint32).reshape(n, 3)


def test_writer_with_device_presort_reproduces_the_reference_test_bin(pkg, tmp_path):
    """flat 128 x 128 grid -> rtrt_write_mesh_bin(presort, legacy layout) must be the reference's resources/models/test.bin
    (positions bit for bit; its pad words are uninitialised memory)."""
    g = mf.flat_grid(128)
    p, i = soup(g)
    path = str(tmp_path / "grid.bin")
    with pkg.RayTracer(64, 64) as rt:
        rt.set_mesh(p, i)
        rt.write_mesh_bin(path, presort=True, legacy64=True)
        assert os.path.getsize(path) == 4 + 64 * 32768                 # the size of the shipped asset
        pos, nrm = mf.read_bin(path)
        assert nrm is None and sha(pos) == GOLD["test.bin"]["sha256_positions"]
        np.testing.assert_array_equal(rt.read_buffer(pkg.BUF_TRI_PERMUTATION), mf.presort_order(g).astype(np.uint32))
        # and the legacy file loads back into a buildable mesh
        rt.set_mesh_bin(path); rt.build_bvh()
        tri = rt.read_buffer(pkg.BUF_TRIANGLES)
    assert tri.shape[0] == 32768


@pytest.mark.parametrize("presort", [False, True])
def test_bin_round_trip_and_order(pkg, orc, tmp_path, presort):
    pos, nrm, idx = pkg.scenes.terrain(40, 52, seed=5, order="shuffle")
    tri_pos, tri_nrm = pos[idx], nrm[idx]
    a, b = str(tmp_path / "a.bin"), str(tmp_path / "b.bin")
    with pkg.RayTracer(64, 64) as rt:
        rt.set_mesh(pos, idx, nrm)
        rt.write_mesh_bin(a, presort=presort)
        order = mf.presort_order(tri_pos) if presort else np.arange(idx.shape[0])
        got_p, got_n = mf.read_bin(a)
        np.testing.assert_array_equal(got_p, tri_pos[order]); np.testing.assert_array_equal(got_n, tri_nrm[order])
        # reading the file back and writing it unsorted again is the identity
        rt.set_mesh_bin(a); rt.write_mesh_bin(b, presort=False)
        assert open(a, "rb").read() == open(b, "rb").read()
        # the mesh built from the file = the oracle's build of the same triangles
        rt.set_option("keep_presort_morton", 1); rt.build_bvh()
        sp, si = soup(tri_pos[order])
        ob = orc.build_bvh(sp, tri_nrm[order].reshape(-1, 3), si)
        np.testing.assert_array_equal(rt.read_buffer(pkg.BUF_MORTON), ob.morton)
        np.testing.assert_array_equal(rt.read_buffer(pkg.BUF_BVH_NODES), ob.nodes)


def test_obj_ingest_builds_like_the_oracle(pkg, orc, tmp_path):
    path = write_obj(tmp_path, OBJ_TEXT)
    opos, onrm = mf.read_obj(path)
    with pkg.RayTracer(64, 64) as rt:
        rt.set_mesh_obj(path); rt.build_bvh()
        sp, si = soup(opos)
        ob = orc.build_bvh(sp, onrm.reshape(-1, 3), si)
        np.testing.assert_array_equal(rt.read_buffer(pkg.BUF_BVH_NODES), ob.nodes)
        np.testing.assert_array_equal(rt.read_buffer(pkg.BUF_TRIANGLES), ob.triangles)
        with pytest.raises(pkg.RtrtError, match="line 4"):
            rt.set_mesh_obj(write_obj(tmp_path, "v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 9\n", name="bad.obj"))
        with pytest.raises(pkg.RtrtError):
            rt.write_mesh_bin(str(tmp_path / "no" / "such" / "dir.bin"))
