#!/bin/bash
# default library: stage timing (primary_persist off) ; variants: parity tests of the path tracer + stage timing
echo "=== default"; NOREF=1 timeout 300 python tools/exp_stages.py kernel_timing=1 2>&1 | tail -1 | cut -c1-400
for v in "$@"; do
  export RTRT_EXPERIMENT=1 RTRT_LIB=$PWD/real-time-ray-tracing_b200/librtrt_$v.so
  echo "=== variant $v"
  timeout 400 python -m pytest tests/test_pathtrace_gpu.py tests/test_refcuda_frame_gpu.py tests/test_parallel_gpu.py -x -q 2>&1 | tail -2
  NOREF=1 timeout 300 python tools/exp_stages.py 2>&1 | tail -1 | cut -c1-400
done
