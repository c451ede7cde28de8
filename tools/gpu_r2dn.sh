#!/bin/bash
# denoiser fast paths: A = in-tree (RTRT_DN_FAST 1, bit-identical claim), dnold = the previous commit's denoise.cu, dnB = explicit FMAs (tolerance only)
OUT=gpurun_out; mkdir -p $OUT
export RTRT_EXPERIMENT=1 RTRT_DATA_DIR=$PWD/real-time-ray-tracing_b200/data
BV=$PWD/build_variants
{
timeout 60 python tools/exp_dn_ab.py
RTRT_LIB=$BV/librtrt_dnold.so timeout 60 python tools/exp_dn_ab.py
RTRT_LIB=$BV/librtrt_dnB.so timeout 60 python tools/exp_dn_ab.py
timeout 40 python tools/exp_dn_ab.py passes=1,1,2,1
RTRT_LIB=$BV/librtrt_dnold.so timeout 40 python tools/exp_dn_ab.py passes=1,1,2,1
} 2>&1 | grep -v Warning | tee $OUT/r2dn_ab.txt
T="tests/test_denoise_gpu.py tests/test_refcuda_frame_gpu.py tests/test_bench_config_gpu.py"
RTRT_LIB=$BV/librtrt_dnB.so timeout 100 python -m pytest $T -x -q -m gpu -k "not c5 and not 64spp and not wavefront" > $OUT/r2dn_pytest_B.log 2>&1; echo "pytest B exit $?"
tail -2 $OUT/r2dn_pytest_B.log
true
