#!/bin/bash
# timing of the given variants, then ncu --set full (source-level) of k_build_blas for the first one
for v in "$@"; do
  echo "=== variant $v"; RTRT_EXPERIMENT=1 RTRT_LIB=$PWD/build_variants/$v.so timeout 200 python tools/exp_build_occ.py 2>&1 | tail -2
done
export RTRT_EXPERIMENT=1 RTRT_LIB=$PWD/build_variants/$1.so
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_build_blas -s 4 -c 1 -o gpurun_out/build_$1 -f python tools/exp_build_occ.py > gpurun_out/build_prof.log 2>&1; echo "ncu exit $?"
