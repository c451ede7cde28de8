#!/bin/bash
# path-trace stage timing of the default library and of the given variants (real-time-ray-tracing_b200/librtrt_<name>.so), 1080p and 4K
run() { NOREF=1 timeout 300 python tools/exp_stages.py kernel_timing=1 2>&1 | tail -1 | grep -o '"W": [0-9]*\|"pathtrace_ms": [0-9.]*\|"denoise_ms": [0-9.]*' | tr '\n' ' '; echo; }
echo "=== default"; run; W=3840 H=2160 run
for v in "$@"; do
  export RTRT_EXPERIMENT=1 RTRT_LIB=$PWD/real-time-ray-tracing_b200/librtrt_$v.so
  echo "=== variant $v"; run; W=3840 H=2160 run
  timeout 300 python -m pytest tests/test_pathtrace_gpu.py tests/test_refcuda_frame_gpu.py -x -q 2>&1 | tail -1
done
