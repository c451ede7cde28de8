#!/bin/bash
# For every build_variants/<name>.so given: BVH parity tests (oracle + reference kernels, bit-exact) and build-kernel timing.
for v in "$@"; do
  export RTRT_EXPERIMENT=1 RTRT_LIB=$PWD/real-time-ray-tracing_b200/librtrt_$v.so
  echo "=== variant $v"
  timeout 300 python -m pytest tests/test_bvh_gpu.py tests/test_refcuda_gpu.py -x -q 2>&1 | tail -2
  timeout 200 python tools/exp_build_occ.py 2>&1 | tail -2
done
