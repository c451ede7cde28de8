"""SURVEY 8f N4 — mesh ingest formats, host side (no GPU): the product's context-free readers (rtrt_mesh_file_read) against the oracle
restatement (oracle/mesh_formats.py), the oracle against the reference's own assets (only where /root/reference exists) and
against the committed digests of those assets (tests/golden/mesh_formats.json, made by tools/make_golden_mesh.py)."""
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import mesh_formats as mf

REF = "/root/reference/resources/models"
GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "mesh_formats.json")))
needs_reference = pytest.mark.skipif(not os.path.isdir(REF), reason="reference assets are only in the build container")


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, np.float32).tobytes()).hexdigest()


OBJ_TEXT = """# comment line
o thing
v 0 0 0
v 1 0 0
v 1 1 0
v 0 1 0   # trailing comment
v 0.5 0.5 1.25e0
vt 0 0
vt 1 1
vn 0 0 1
vn 0 1 0
vn 1 0 0
f 1/1/1 2/2/1 3/1/1
f 1//1 3//1 4//1
f -5/1/2 -4/2/2 -1/1/3
f 1/1/1 2/1/2 3/1/3 4/1/1
f 2/1/1 3/1/1 5/1/2 4/1/3 1/1/1
usemtl whatever
s off
"""


def write_obj(tmp_path, text=OBJ_TEXT, name="t.obj", crlf=False):
    p = tmp_path / name
    p.write_bytes(text.replace("\n", "\r\n").encode() if crlf else text.encode())
    return str(p)


@pytest.mark.parametrize("crlf", [False, True])
def test_obj_reader_matches_oracle(pkg, tmp_path, crlf):
    path = write_obj(tmp_path, crlf=crlf)
    pos, nrm = pkg.read_mesh_file(path)
    opos, onrm = mf.read_obj(path)
    assert pos.shape == (1 + 1 + 1 + 2 + 3, 3, 3)                    # triangle, triangle, relative, quad fan, pentagon fan
    np.testing.assert_array_equal(pos, opos); np.testing.assert_array_equal(nrm, onrm)
    np.testing.assert_array_equal(pos[2], np.array([[0, 0, 0], [1, 0, 0], [0.5, 0.5, 1.25]], np.float32))      # negative indices
    np.testing.assert_array_equal(pos[3:5, 0], np.zeros((2, 3), np.float32))                                    # fan from corner 0
    np.testing.assert_array_equal(nrm[2], np.array([[0, 1, 0], [0, 1, 0], [1, 0, 0]], np.float32))


def test_obj_without_normals_reports_none(pkg, tmp_path):
    path = write_obj(tmp_path, "v 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nf 1//1 2//1 3//1\nf 1 2 3\n")
    pos, nrm = pkg.read_mesh_file(path)
    assert pos.shape == (2, 3, 3) and nrm is None and mf.read_obj(path)[1] is None


@pytest.mark.parametrize("text,what", [("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 4\n", "out of range"), ("v 0 0 0\nv 1 0\n", "malformed vertex"),
                                       ("v 0 0 0\nv 1 0 0\nf 1 2\n", "fewer than 3"), ("v 0 0 0\n", "no faces"),
                                       ("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 x 3\n", "malformed face")])
def test_obj_errors_are_reported_with_context(pkg, tmp_path, text, what):
    path = write_obj(tmp_path, text)
    with pytest.raises(pkg.RtrtError, match=what):
        pkg.read_mesh_file(path)


def test_missing_file(pkg, tmp_path):
    for name in ("nope.obj", "nope.bin"):
        with pytest.raises(pkg.RtrtError, match="cannot open"):
            pkg.read_mesh_file(str(tmp_path / name))


@pytest.mark.parametrize("legacy", [False, True])
def test_bin_reader_matches_oracle_both_layouts(pkg, tmp_path, legacy):
    rng = np.random.default_rng(3)
    pos = rng.normal(size=(1000, 3, 3)).astype(np.float32); nrm = rng.normal(size=(1000, 3, 3)).astype(np.float32)
    path = str(tmp_path / "m.bin")
    mf.write_bin(path, pos, nrm, legacy64=legacy)
    assert os.path.getsize(path) == 4 + (64 if legacy else 128) * 1000
    got_p, got_n = pkg.read_mesh_file(path)
    np.testing.assert_array_equal(got_p, pos)
    if legacy:
        assert got_n is None
    else:
        np.testing.assert_array_equal(got_n, nrm)
    with open(path, "r+b") as f:                                       # truncated payload
        f.truncate(4 + 50 * 1000)
    with pytest.raises(pkg.RtrtError, match="truncated"):
        pkg.read_mesh_file(path)
    with open(path, "wb") as f:
        f.write(b"\0\0")
    with pytest.raises(pkg.RtrtError, match="bad header"):
        pkg.read_mesh_file(path)


def test_flat_grid_presorted_reproduces_the_reference_test_bin_digest():
    """resources/models/test.bin is the output of the reference's mesh tool on a 128 x 128 grid: the oracle's restatement of the tool
    (60-bit Morton key of the centroid, sort, write) applied to the same grid reproduces the asset's triangles bit for bit."""
    g = mf.flat_grid(128)
    assert g.shape[0] == GOLD["test.bin"]["triangles"] == 32768
    assert sha(g[mf.presort_order(g)]) == GOLD["test.bin"]["sha256_positions"]
    keys = mf.morton60_keys(g)
    assert np.unique(keys).size == keys.size                           # no ties: the order is fully determined


@needs_reference
def test_golden_digests_are_those_of_the_reference_assets(pkg):
    pos, nrm = mf.read_bin(os.path.join(REF, "test.bin"))
    assert nrm is None and sha(pos) == GOLD["test.bin"]["sha256_positions"]
    ppos, pnrm = pkg.read_mesh_file(os.path.join(REF, "test.bin"))      # the product reads the legacy asset too
    assert pnrm is None and sha(ppos) == GOLD["test.bin"]["sha256_positions"]
    pos, nrm = mf.read_obj(os.path.join(REF, "monkey.obj"))
    assert pos.shape[0] == GOLD["monkey.obj"]["triangles"] == 968       # 968 triangles, 507 vertices
    assert sum(1 for l in open(os.path.join(REF, "monkey.obj")) if l.startswith("v ")) == 507
    assert np.unique(pos.reshape(-1, 3), axis=0).shape[0] == GOLD["monkey.obj"]["unique_vertices"]      # 505: two `v` records coincide
    assert sha(pos) == GOLD["monkey.obj"]["sha256_positions"] and sha(nrm) == GOLD["monkey.obj"]["sha256_normals"]
    ppos, pnrm = pkg.read_mesh_file(os.path.join(REF, "monkey.obj"))
    assert sha(ppos) == GOLD["monkey.obj"]["sha256_positions"] and sha(pnrm) == GOLD["monkey.obj"]["sha256_normals"]


@needs_reference
def test_product_reads_every_roundcubes_obj_like_the_oracle(pkg):
    import glob
    files = sorted(glob.glob(os.path.join(REF, "roundcubes", "*", "*.obj")))
    assert len(files) == GOLD["roundcubes"]["files"]
    h = hashlib.sha256(); total = 0
    for f in files:
        pos, nrm = pkg.read_mesh_file(f)
        opos, onrm = mf.read_obj(f)
        np.testing.assert_array_equal(pos, opos)
        assert (nrm is None) == (onrm is None) and (nrm is None or np.array_equal(nrm, onrm))
        h.update(np.ascontiguousarray(pos).tobytes()); total += pos.shape[0]
    assert total == GOLD["roundcubes"]["triangles"] and h.hexdigest() == GOLD["roundcubes"]["sha256_positions_concatenated"]
