#!/usr/bin/env python
"""Writes tests/golden/mesh_formats.json: digests of the reference's own mesh assets as read by the oracle restatement
(oracle/mesh_formats.py).  Run in the container that has /root/reference; the GPU box has only the JSON.

    python tools/make_golden_mesh.py
"""
import glob
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from oracle import mesh_formats as mf  # noqa: E402

REF = "/root/reference/resources/models"


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, np.float32).tobytes()).hexdigest()


def main():
    out = {"_made_by": "tools/make_golden_mesh.py", "_source": "wangkepfe/Real-Time-Ray-Tracing resources/models"}
    pos, nrm = mf.read_bin(os.path.join(REF, "test.bin"))
    assert nrm is None
    out["test.bin"] = {"layout": "legacy64", "triangles": int(pos.shape[0]), "sha256_positions": sha(pos),
                       "bounds": [pos.reshape(-1, 3).min(axis=0).tolist(), pos.reshape(-1, 3).max(axis=0).tolist()]}
    pos, nrm = mf.read_obj(os.path.join(REF, "monkey.obj"))
    out["monkey.obj"] = {"triangles": int(pos.shape[0]), "unique_vertices": int(np.unique(pos.reshape(-1, 3), axis=0).shape[0]),
                         "sha256_positions": sha(pos), "sha256_normals": sha(nrm),
                         "sha256_positions_presorted": sha(pos[mf.presort_order(pos)])}
    files = sorted(glob.glob(os.path.join(REF, "roundcubes", "*", "*.obj")))
    h = hashlib.sha256(); total = 0
    for f in files:
        p, n = mf.read_obj(f)
        h.update(np.ascontiguousarray(p, np.float32).tobytes()); total += int(p.shape[0])
    out["roundcubes"] = {"files": len(files), "triangles": total, "sha256_positions_concatenated": h.hexdigest()}
    path = os.path.join(ROOT, "tests", "golden", "mesh_formats.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print(path, json.dumps(out)[:300])


if __name__ == "__main__":
    main()
