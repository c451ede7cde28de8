#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
timeout 300 python -m pytest tests/test_multi_gpu.py -x -q -m gpu 2>&1 | tail -15
timeout 1500 python -m pytest tests -x -q -m gpu --durations=8 > $OUT/r2g_pytest.log 2>&1; echo "pytest exit $?"
tail -14 $OUT/r2g_pytest.log
timeout 200 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -2
python tools/exp_wide_ab.py wide_bvh=1
