#!/bin/bash
export RTRT_DATA_DIR=/root/repo/real-time-ray-tracing_b200/data
for v in r4 r8 t2 t5; do RTRT_EXPERIMENT=1 RTRT_LIB=/root/repo/build_variants/librtrt_$v.so python tools/exp_wide_ab.py wide_bvh=1; done
python tools/exp_wide_ab.py wide_bvh=1 > /dev/null 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_pt_trace|k_pt_primary' -s 4 -c 3 -o gpurun_out/r2d_prof -f python tools/exp_wide_ab.py wide_bvh=1 > gpurun_out/r2d_ncu.log 2>&1
echo ncu exit $?
