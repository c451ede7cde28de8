#!/bin/bash
export RTRT_DATA_DIR=/root/repo/real-time-ray-tracing_b200/data
python tools/exp_wide_ab.py wide_bvh=1
for v in p2 p3 p4; do RTRT_EXPERIMENT=1 RTRT_LIB=/root/repo/build_variants/librtrt_$v.so python tools/exp_wide_ab.py wide_bvh=1 2>&1 | tail -1; done
