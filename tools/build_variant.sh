#!/bin/bash
# usage: tools/build_variant.sh <name> "<extra nvcc flags>"   -> build_variants/librtrt_<name>.so   (A/B experiments, RTRT_EXPERIMENT=1 RTRT_LIB=...)
mkdir -p /root/repo/build_variants
RTRT_BUILD_OUT=/root/repo/build_variants/librtrt_$1.so RTRT_NVCC_EXTRA="$2" python /root/repo/real-time-ray-tracing_b200/build.py --force > /tmp/build_$1.log 2>&1 && echo "built $1" || { echo "FAILED $1"; grep -i error /tmp/build_$1.log | head -5; }
