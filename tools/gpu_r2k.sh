#!/bin/bash
python -m pytest tests/test_wide_bvh_gpu.py tests/test_multi_gpu.py tests/test_async_gpu.py -x -q -m gpu 2>&1 | tail -3
python tools/exp_wide_ab.py wide_bvh=1
python tools/exp_wide_ab.py wide_bvh=2
python tools/exp_wide_ab.py wide_bvh=0
