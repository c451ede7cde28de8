#!/bin/bash
# path-trace stage timing of the default library and of the given variants (real-time-ray-tracing_b200/librtrt_<name>.so)
echo "=== default"; NOREF=1 timeout 300 python tools/exp_stages.py 2>&1 | tail -1 | cut -c70-330
for v in "$@"; do
  echo "=== variant $v"; RTRT_LIB=$PWD/real-time-ray-tracing_b200/librtrt_$v.so NOREF=1 timeout 300 python tools/exp_stages.py 2>&1 | tail -1 | cut -c70-330
done
