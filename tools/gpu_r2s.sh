#!/bin/bash
python -m pytest tests/test_pathtrace_gpu.py tests/test_traverse_gpu.py tests/test_bench_config_gpu.py -x -q -m gpu 2>&1 | tail -2
python tools/exp_wide_ab.py wide_bvh=0
python tools/exp_wide_ab.py wide_bvh=1
