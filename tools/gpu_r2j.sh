#!/bin/bash
python -m pytest tests/test_wide_bvh_gpu.py tests/test_bench_config_gpu.py -x -q -m gpu 2>&1 | tail -3
python tools/exp_wide_ab.py wide_bvh=1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_wide_|k_build' -c 12 python tools/exp_wide_ab.py wide_bvh=1 2>&1 | grep -E "k_wide|k_build|gpu__time" | head -24
