#!/bin/bash
# A/B builds of librtrt.so for one gpurun call: tools/build_variants.sh name1 "-DX=1 -DY=0" name2 "..." -> build_variants/<name>.so
# (git-ignored, travels to the GPU box; select with RTRT_LIB=build_variants/<name>.so).  The default library is rebuilt at the end.
set -e
cd "$(dirname "$0")/.."
mkdir -p build_variants
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  RTRT_NVCC_EXTRA="$flags" python real-time-ray-tracing_b200/build.py --force > /dev/null
  cp real-time-ray-tracing_b200/librtrt.so build_variants/$name.so; cp real-time-ray-tracing_b200/librtrt.so real-time-ray-tracing_b200/librtrt_$name.so
  echo "built build_variants/$name.so  [$flags]"
done
python real-time-ray-tracing_b200/build.py --force > /dev/null
