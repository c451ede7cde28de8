#!/bin/bash
export RTRT_DATA_DIR=/root/repo/real-time-ray-tracing_b200/data
python -m pytest tests/test_wide_bvh_gpu.py -x -q -m gpu 2>&1 | tail -2
python tools/wide_stats.py | grep triangles
python tools/exp_wide_ab.py wide_bvh=1
RTRT_EXPERIMENT=1 RTRT_LIB=/root/repo/build_variants/librtrt_nodis.so python tools/exp_wide_ab.py wide_bvh=1 2>&1 | tail -1
