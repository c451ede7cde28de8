#!/bin/bash
python -m pytest tests/test_wide_bvh_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu -s 2>&1 | grep -E "passed|failed|differ|Error|assert" | head -20
python tools/exp_wide_ab.py wide_bvh=0
python tools/exp_wide_ab.py wide_bvh=1
for v in r4 r8 t2 t5; do RTRT_EXPERIMENT=1 RTRT_LIB=/root/repo/build_variants/librtrt_$v.so python tools/exp_wide_ab.py wide_bvh=1; done
