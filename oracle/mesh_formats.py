"""ORACLE — TEST INFRASTRUCTURE ONLY (the on-disk mesh formats either side of the hot path).

Plain Python / numpy restatement of
  * `.bin` reader  LoadTrianglesFromFile            src/init.cu:28-50      uint32 triCount; Triangle[triCount]
  * `.bin` writer + offline 60-bit Morton pre-sort   tool/meshProcessor.cpp:35-62 (code), :146-171 (per-triangle key), :174-208 (main)
  * Triangle record                                  src/geometry.h:52-110  v1 w1 v2 w2 v3 w3 v4 w4 n1 u1 n2 u2 n3 u3 n4 u4 (128 B);
                                                     the shipped resources/models/test.bin uses the older 64-B record (v1..v4 only)
  * OBJ ingest      LoadScene / LoadSceneRecursive   src/fileUtils.cu:16-80 — the reference goes through Assimp 5.x (un-vendored
                    dependency, `aiProcess_Triangulate`); restated here is the published OBJ grammar subset the reference's assets use
                    (v, vn, f with v | v/vt | v//vn | v/vt/vn corners, negative indices) and Assimp's output order for a single-object
                    file: faces in file order, polygons emitted as a fan from the first corner.

Parity status: pinned on the reference's own assets when /root/reference is present (tests/test_mesh_formats_cpu.py:
resources/models/monkey.obj = 968 triangles / 507 vertices, resources/models/test.bin = 32 768 legacy triangles of a 128 x 128 grid);
their digests are committed in tests/golden/mesh_formats.json (made by tools/make_golden_mesh.py) so the GPU box checks against them.
Only tests/ may import this module.
"""
import numpy as np

TRI128 = np.dtype([("v", "<f4", (4, 4)), ("n", "<f4", (4, 4))])        # rows: xyz + pad word
TRI64 = np.dtype([("v", "<f4", (4, 4))])


def read_bin(path):
    """-> (pos [n,3,3] f32, nrm [n,3,3] f32 or None).  The record size follows from the file size (both layouts share the header)."""
    raw = open(path, "rb").read()
    n = int(np.frombuffer(raw[:4], "<u4")[0])
    if len(raw) == 4 + 64 * n:
        t = np.frombuffer(raw, TRI64, count=n, offset=4)
        return t["v"][:, :3, :3].copy(), None
    if len(raw) < 4 + 128 * n:
        raise ValueError("truncated .bin")
    t = np.frombuffer(raw, TRI128, count=n, offset=4)
    return t["v"][:, :3, :3].copy(), t["n"][:, :3, :3].copy()


def write_bin(path, pos, nrm=None, legacy64=False):
    """tool/meshProcessor.cpp:201-208.  Pad words, v4 and n4 are written as zero."""
    pos = np.asarray(pos, np.float32).reshape(-1, 3, 3)
    n = pos.shape[0]
    t = np.zeros(n, TRI64 if legacy64 else TRI128)
    t["v"][:, :3, :3] = pos
    if not legacy64:
        t["n"][:, :3, :3] = np.asarray(nrm, np.float32).reshape(-1, 3, 3)
    with open(path, "wb") as f:
        f.write(np.uint32(n).tobytes()); f.write(t.tobytes())


def _expand10(x):
    """MortonCode3D's bit spreading, tool/meshProcessor.cpp:35-40 (10 bits -> every third bit)."""
    x = x.astype(np.uint32)
    x = (x | (x << np.uint32(16))) & np.uint32(0x030000FF)
    x = (x | (x << np.uint32(8))) & np.uint32(0x0300F00F)
    x = (x | (x << np.uint32(4))) & np.uint32(0x030C30C3)
    x = (x | (x << np.uint32(2))) & np.uint32(0x09249249)
    return x


def morton60_keys(pos):
    """CalculateAabb + CalculateMortonOrder, tool/meshProcessor.cpp:121-171, float32 like the tool:
    centre = (v1 + v2 + v3) / 3.0f; unitBox = (centre - sceneMin) / (sceneMax - sceneMin); q = (uint)(unitBox * 1048575.f);
    key = Morton30(q >> 10) << 30 | Morton30(q & 1023).  A flat axis (max == min) is 0/0 in the tool (undefined conversion);
    it is defined as q = 0 here and in the product."""
    f32 = np.float32
    pos = np.asarray(pos, f32).reshape(-1, 3, 3)
    lo = pos.reshape(-1, 3).min(axis=0); hi = pos.reshape(-1, 3).max(axis=0)
    centre = ((pos[:, 0] + pos[:, 1]) + pos[:, 2]) / f32(3.0)
    ext = (hi - lo).astype(f32)
    ext = np.where(ext > 0, ext, f32(1.0)).astype(f32)
    unit = ((centre - lo) / ext).astype(f32)
    q = np.clip(unit * f32(1048575.0), f32(0), f32(1048575.0)).astype(np.uint32)      # (uint) truncates
    def m30(a):
        return (_expand10(a[:, 0]) | (_expand10(a[:, 1]) << np.uint32(1)) | (_expand10(a[:, 2]) << np.uint32(2))).astype(np.uint64)
    return (m30(q >> np.uint32(10)) << np.uint64(30)) | m30(q & np.uint32(1023))


def presort_order(pos):
    """main(), tool/meshProcessor.cpp:188-199: triangle order of the written file.  The tool's std::sort leaves equal keys in
    unspecified order; the stable order (ties keep input order) is the one the product promises."""
    return np.argsort(morton60_keys(pos), kind="stable")


def read_obj(path):
    """-> (pos [n,3,3] f32, nrm [n,3,3] f32 or None if any face corner lacks a normal — aiMesh::HasNormals is per mesh)."""
    v, vn, tp, tn = [], [], [], []
    all_vn = True
    with open(path, "r", errors="replace") as f:
        for line_no, line in enumerate(f, start=1):
            tok = line.split("#", 1)[0].split()
            if not tok:
                continue
            if tok[0] == "v":
                v.append([np.float32(t) for t in tok[1:4]])
            elif tok[0] == "vn":
                vn.append([np.float32(t) for t in tok[1:4]])
            elif tok[0] == "f":
                cv, cn = [], []
                for c in tok[1:]:
                    parts = c.split("/")
                    a = int(parts[0]); a = len(v) + a if a < 0 else a - 1
                    if not 0 <= a < len(v):
                        raise ValueError(f"vertex index out of range at line {line_no}")
                    b = None
                    if len(parts) == 3 and parts[2] != "":
                        b = int(parts[2]); b = len(vn) + b if b < 0 else b - 1
                        if not 0 <= b < len(vn):
                            raise ValueError(f"normal index out of range at line {line_no}")
                    cv.append(a); cn.append(b)
                if len(cv) < 3:
                    raise ValueError(f"face with fewer than 3 corners at line {line_no}")
                face_vn = all(b is not None for b in cn)
                all_vn = all_vn and face_vn
                for k in range(1, len(cv) - 1):
                    tp.append([v[cv[0]], v[cv[k]], v[cv[k + 1]]])
                    tn.append([vn[cn[0]], vn[cn[k]], vn[cn[k + 1]]] if face_vn else [[0, 0, 0]] * 3)
    if not tp:
        raise ValueError("no faces in file")
    return np.asarray(tp, np.float32), (np.asarray(tn, np.float32) if all_vn else None)


def flat_grid(cells=128):
    """The mesh of the reference's resources/models/test.bin: a cells x cells grid of quads on z = 0 over [-1,1]^2, two triangles per
    quad, 32 768 triangles for cells = 128.  Order and winding reproduce the asset (checked against it in tests/test_mesh_formats_cpu.py)."""
    f32 = np.float32
    step = f32(2.0) / f32(cells)
    i, j = np.meshgrid(np.arange(cells), np.arange(cells), indexing="xy")
    i = i.reshape(-1).astype(f32); j = j.reshape(-1).astype(f32)
    x0, y0 = f32(-1.0) + i * step, f32(-1.0) + j * step
    x1, y1 = x0 + step, y0 + step
    z = np.zeros_like(x0)
    P = lambda x, y: np.stack([x, y, z], axis=-1)
    t0 = np.stack([P(x1, y0), P(x0, y1), P(x0, y0)], axis=1)
    t1 = np.stack([P(x1, y0), P(x1, y1), P(x0, y1)], axis=1)
    return np.stack([t0, t1], axis=1).reshape(-1, 3, 3).astype(np.float32)
