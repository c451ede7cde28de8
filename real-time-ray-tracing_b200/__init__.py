"""real-time-ray-tracing_b200 — B200-native (sm_100a) hot path of wangkepfe/Real-Time-Ray-Tracing:
LBVH build -> BVH traversal / path trace -> temporal + a-trous denoise, behind a C ABI (include/rtrt.h).

The directory name is not a Python identifier; import it with
    importlib.import_module("real-time-ray-tracing_b200")
(tests/conftest.py, bench.py and __graft_entry__.py do exactly that).
"""
from . import _lib  # noqa: F401
from .api import (RayTracer, RtrtError, load_config, config_defaults, TRIANGLE_DTYPE, AABB_DTYPE, BVHNODE_DTYPE,
                  RAY_DTYPE, HIT_DTYPE, RAY_MAX, MATERIAL_DTYPE, COLOR_DTYPE, dynamic_resolution_step, read_mesh_file)  # noqa: F401
from ._lib import *  # noqa: F401,F403  (BUF_* ids, ctypes structs)
from . import scenes  # noqa: F401
from . import parallel  # noqa: F401
from .parallel import RowShardedRayTracer, row_bands  # noqa: F401
from ._lib import (BUF_TRIANGLES, BUF_AABBS, BUF_MORTON_PRESORT, BUF_MORTON, BUF_REORDER_IDX, BUF_BVH_NODES,
                   BUF_TLAS_AABBS, BUF_TLAS_MORTON_PRESORT, BUF_TLAS_MORTON, BUF_TLAS_REORDER_IDX,
                   BUF_TLAS_BVH_NODES, BUF_TRI_PERMUTATION)  # noqa: F401
